// Device-resident matcher (include/os1_b200.h, "os1_matcher"): the three matching functions named by the north star as
// WHOLE functions on the GPU — candidate generation, distances, the ordered claim pass, the rotation filter and the
// result arrays — with no host round trip in between and no allocation in a steady-state call.
//
//   ORBmatcher::SearchForInitialization                         src/ORBmatcher.cc:400-515
//   ORBmatcher::SearchByProjection(Frame&, const Frame&, th)    src/ORBmatcher.cc:1292-1423   (incl. the projection :1304-1347)
//   ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&)  src/ORBmatcher.cc:45-124
//
// Pipeline of a call (all on the caller's stream):
//   1. grid_build_kernel                 Frame::AssignFeaturesToGrid of the searched frame              (framegrid.cu)
//   2. a query kernel                    what each query hands to GetFeaturesInArea, or "skipped"        (here)
//   3. area count / scan / fill          Frame::GetFeaturesInArea for every query -> CSR lists           (framegrid.cu)
//   4. hamming_csr_kernel                DescriptorDistance of every (query, candidate)                  (hamming.cu)
//   5. a claim kernel, ONE warp          the reference's in-order loop: a query's outcome depends on what earlier queries
//                                        claimed (:439, :458-466, :1363-1381, :83-85, :116), so queries are taken one after the
//                                        other; inside a query the lanes share the candidate list and reduce (best, second) as
//                                        order statistics of dist << 20 | position keys, which reproduces the scan's strict '<'
//                                        tie rule exactly.  Claim state lives in shared memory; the next query's candidates are
//                                        fetched while the current one is reduced.
//   6. a finish kernel                   ComputeThreeMaxima + rotation filter (:468-505, :1384-1419), vbPrevMatched (:507-510)
#include "../../include/os1_b200.h"
#include "os1_internal.cuh"
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <string>

#define MD_POSBITS 20
#define MD_POSMASK ((1u << MD_POSBITS) - 1)
#define MD_EMPTY 0xFFFFFFFFu
#define MD_HISTO 30            // HISTO_LENGTH  ORBmatcher.cc:39
#define MD_TH_LOW 50           // ORBmatcher.cc:38
#define MD_TH_HIGH 100         // ORBmatcher.cc:37

struct os1_matcher {
    int device = 0, maxKp = 0, maxQ = 0, maxCand = 0;
    // device scratch, sized at creation
    int *cellStart = nullptr, *gridIdx = nullptr, *cnt = nullptr, *ptr = nullptr, *cand = nullptr;
    uint16_t* dist = nullptr;
    float *qx = nullptr, *qy = nullptr, *qr = nullptr;
    int *qmin = nullptr, *qmax = nullptr;
    uint8_t* qvalid = nullptr;
    int *binOf = nullptr, *idxOf = nullptr, *hist = nullptr, *status = nullptr;
    // staging for the host-buffer convenience path (grown on demand, kept)
    uint8_t* stage = nullptr; size_t stageBytes = 0;
    int* hostStatus = nullptr;
    std::string err;
};

namespace {

thread_local std::string t_merr;
int mfail(os1_matcher* m, int code, const std::string& msg)
{
    if (m) m->err = msg;
    os1_internal_set_error(msg.c_str());
    return code;
}
#define MCK(call, what)                                                                                        \
    do {                                                                                                       \
        cudaError_t _e = (call);                                                                               \
        if (_e != cudaSuccess) { cudaGetLastError(); return mfail(m, OS1_E_CUDA, std::string(what) + ": " + cudaGetErrorString(_e)); } \
    } while (0)

// ---- 2. query kernels --------------------------------------------------------------------------------------------------
// SearchForInitialization :413-420: level-0 keypoints of F1 query a window around vbPrevMatched[i1], levels [level1, level1]
__global__ void init_query_kernel(const os1_kp_dev* kp1, int n1, const float* prev, float window, float* qx, float* qy, float* qr, int* qmin, int* qmax,
                                  uint8_t* qvalid)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n1) return;
    const int level1 = kp1[i].octave;
    qvalid[i] = level1 > 0 ? 0 : 1;
    qx[i] = prev[2 * i]; qy[i] = prev[2 * i + 1]; qr[i] = window; qmin[i] = level1; qmax[i] = level1;
}

struct ProjParams { float T[12]; float fx, fy, cx, cy; float minX, maxX, minY, maxY; float th; float scale[OS1_MAX_LEVELS]; int nlevels; };

// SearchByProjection(Frame&, const Frame&, th) :1316-1349.  `Rcw*x3Dw+tcw` is a cv::gemm of a 3x3 by a 3x1 float matrix: float
// products and float sums from left to right, then (float)(t*alpha + c*beta) in double (gemm's small-matrix path; pinned against
// cv2.gemm by tests/test_matcher_reference.py).  Every float operation is spelled with its rounding so nothing contracts.
__global__ void project_query_kernel(const os1_kp_dev* kpl, int nl, const uint8_t* has, const uint8_t* outlier, const float* wpos, ProjParams P,
                                     float* qx, float* qy, float* qr, int* qmin, int* qmax, uint8_t* qvalid)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nl) return;
    qvalid[i] = 0; qx[i] = 0.f; qy[i] = 0.f; qr[i] = 0.f; qmin[i] = -1; qmax[i] = -1;
    if (!has[i] || outlier[i]) return;                                       // :1316, :1320
    const float X = wpos[3 * i], Y = wpos[3 * i + 1], Z = wpos[3 * i + 2];
    float c[3];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        float t = __fmul_rn(P.T[4 * r], X);
        t = __fadd_rn(t, __fmul_rn(P.T[4 * r + 1], Y));
        t = __fadd_rn(t, __fmul_rn(P.T[4 * r + 2], Z));
        c[r] = __double2float_rn(__dadd_rn(__dmul_rn((double)t, 1.0), __dmul_rn((double)P.T[4 * r + 3], 1.0)));
    }
    const float invzc = __double2float_rn(__ddiv_rn(1.0, (double)c[2]));     // :1329
    if (invzc < 0) return;                                                   // :1332
    const float u = __fadd_rn(__fmul_rn(__fmul_rn(P.fx, c[0]), invzc), P.cx);   // :1335
    const float v = __fadd_rn(__fmul_rn(__fmul_rn(P.fy, c[1]), invzc), P.cy);
    if (u < P.minX || u > P.maxX) return;                                    // :1339
    if (v < P.minY || v > P.maxY) return;
    const int o = kpl[i].octave;                                             // LastFrame.mvKeys[i].octave :1344
    if (o < 0 || o >= P.nlevels) return;
    qx[i] = u; qy[i] = v; qr[i] = __fmul_rn(P.th, P.scale[o]); qmin[i] = o - 1; qmax[i] = o + 1; qvalid[i] = 1;
}

// SearchByProjection(Frame&, const vector<MapPoint*>&, th) :51-71.  mp rows: mTrackProjX, mTrackProjY, mTrackViewCos,
// mnTrackScaleLevel, flags (1 mbTrackInView, 2 isBad, 4 plCandidato), unused.
__global__ void point_query_kernel(const float* mp, int n, ProjParams P, float* qx, float* qy, float* qr, int* qmin, int* qmax, uint8_t* qvalid)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int flags = (int)mp[6 * i + 4];
    const int level = (int)mp[6 * i + 3];
    qvalid[i] = 0; qx[i] = 0.f; qy[i] = 0.f; qr[i] = 0.f; qmin[i] = -1; qmax[i] = -1;
    if (!(flags & 1) || (flags & 2) || level < 0 || level >= P.nlevels) return;
    float r = (flags & 4) ? 4.0f : ((double)mp[6 * i + 2] > 0.998 ? 2.5f : 4.0f);       // RadiusByViewingCos :126-132
    if ((double)P.th != 1.0) r = __fmul_rn(r, P.th);                                     // bFactor :49, :67
    qx[i] = mp[6 * i]; qy[i] = mp[6 * i + 1]; qr[i] = __fmul_rn(r, P.scale[level]); qmin[i] = level - 1; qmax[i] = level; qvalid[i] = 1;
}

// ---- 5. claim kernels: one warp ------------------------------------------------------------------------------------------
__device__ __forceinline__ void warp_merge(unsigned& best, unsigned& second)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned ob = __shfl_xor_sync(0xffffffffu, best, o), os = __shfl_xor_sync(0xffffffffu, second, o);
        second = min(min(second, os), max(best, ob));
        best = min(best, ob);
    }
}

__device__ __forceinline__ int rot_bin(float a1, float a2)
{
    float rot = __fsub_rn(a1, a2);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    const float factor = 1.0f / MD_HISTO;
    int bin = (int)roundf(__fmul_rn(rot, factor));
    if (bin == MD_HISTO) bin = 0;
    return bin;
}

// MODE 0: SearchForInitialization (state: distance each F2 keypoint was claimed with, and by whom)
// MODE 1: SearchByProjection(Frame, Frame)   (state: taken flags; best only)
// MODE 2: SearchByProjection(Frame, points)  (state: taken flags; best + second with their levels)
struct ClaimParams {
    int nq, nc;
    const int* ptr; const int* cand; const uint16_t* dist;
    int candCap;
    float nnratio;
    int checkOri;
    const os1_kp_dev* kpq;      // query-side keypoints (angles)           MODE 0, 1
    const os1_kp_dev* kpc;      // searched frame's keypoints (angles, octaves)
    uint8_t* taken;             // MODE 1, 2: in/out, one per searched keypoint (0 none, 1 point with observations, 2 point without)
    const uint8_t* qhas;        // MODE 1: has_point of the queries (2 = the point has no observations: it does not block later queries)
    int* out;                   // MODE 0: matches12[nq];  MODE 1, 2: match_of_cur[nc]
    int* nmatches;
    int* binOf;                 // per query: rotation bin of its accepted match, or -1
    int* idxOf;                 // MODE 1, per query: the searched keypoint it was attached to (the entry rotHist keeps, :1394)
    int* hist;                  // MD_HISTO counters
    int* status;
};

template <int MODE>
__global__ void __launch_bounds__(32) claim_kernel(ClaimParams P)
{
    extern __shared__ unsigned char smraw[];
    // MODE 0: u16 claimed distance (0xFFFF = INT_MAX) + i16 claimant per searched keypoint; MODE 1, 2: u8 taken
    unsigned short* claimedDist = (unsigned short*)smraw;
    short* claimant = (short*)(smraw + 2 * (size_t)P.nc);
    unsigned char* taken = smraw;
    const int lane = threadIdx.x;
    const int total = P.ptr[P.nq];
    if (total > P.candCap) {        // the candidate lists did not fit: nothing below would be exact
        if (lane == 0) { atomicOr(P.status, 1); *P.nmatches = 0; }
        return;
    }
    if (MODE == 0) {
        for (int i = lane; i < P.nc; i += 32) { claimedDist[i] = 0xFFFF; claimant[i] = -1; }
        for (int i = lane; i < P.nq; i += 32) { P.out[i] = -1; P.binOf[i] = -1; }
    } else {
        for (int i = lane; i < P.nc; i += 32) { taken[i] = P.taken[i]; P.out[i] = -1; }
        if (MODE == 1)
            for (int i = lane; i < P.nq; i += 32) P.binOf[i] = -1;
    }
    if (lane < MD_HISTO) P.hist[lane] = 0;
    __syncwarp();
    int nm = 0;
    // candidates of the next query, fetched while the current one is reduced
    int p0n = P.nq > 0 ? P.ptr[0] : 0, p1n = P.nq > 0 ? P.ptr[1] : 0;
    int dn = 0, in_ = 0;
    if (p0n + lane < p1n) { dn = P.dist[p0n + lane]; in_ = P.cand[p0n + lane]; }
    for (int q = 0; q < P.nq; q++) {
        const int p0 = p0n, p1 = p1n;
        int d = dn, i2 = in_;
        if (q + 1 < P.nq) {
            p0n = p1; p1n = P.ptr[q + 2];
            if (p0n + lane < p1n) { dn = P.dist[p0n + lane]; in_ = P.cand[p0n + lane]; }
        }
        if (p0 == p1) continue;                                                  // :417-423 / :1351 / :75
        unsigned best = MD_EMPTY, second = MD_EMPTY;
        for (int k = p0 + lane; k < p1; k += 32) {
            if (k >= p0 + 32) { d = P.dist[k]; i2 = P.cand[k]; }
            bool skip;
            if (MODE == 0) skip = (int)claimedDist[i2] <= d;                     // :439 (0xFFFF: never claimed)
            else skip = taken[i2] == 1;                                          // :1363-1365 / :83-85
            const unsigned key = skip ? MD_EMPTY : (((unsigned)d << MD_POSBITS) | (unsigned)(k - p0));
            second = min(second, max(key, best));
            best = min(best, key);
        }
        warp_merge(best, second);
        if (MODE == 0) {
            const int bestDist = best == MD_EMPTY ? INT_MAX : (int)(best >> MD_POSBITS);
            const int bestDist2 = second == MD_EMPTY ? INT_MAX : (int)(second >> MD_POSBITS);
            if (bestDist <= MD_TH_LOW && (float)bestDist < __fmul_rn((float)bestDist2, P.nnratio)) {      // :454-456
                const int bestIdx2 = P.cand[p0 + (int)(best & MD_POSMASK)];
                if (lane == 0) {
                    const int old = claimant[bestIdx2];
                    if (old >= 0) { P.out[old] = -1; nm--; }                     // :458-462
                    P.out[q] = bestIdx2; claimant[bestIdx2] = (short)q; claimedDist[bestIdx2] = (unsigned short)bestDist; nm++;
                    if (P.checkOri) {
                        const int bin = rot_bin(P.kpq[q].angle, P.kpc[bestIdx2].angle);      // :468-478
                        P.binOf[q] = bin; P.hist[bin]++;
                    }
                }
                __syncwarp();
            }
        } else if (MODE == 1) {
            const int bestDist = best == MD_EMPTY ? 256 : (int)(best >> MD_POSBITS);
            if (bestDist <= MD_TH_HIGH) {                                        // :1379
                const int bestIdx2 = P.cand[p0 + (int)(best & MD_POSMASK)];
                if (lane == 0) {
                    P.out[bestIdx2] = q; taken[bestIdx2] = P.qhas[q] == 2 ? 2 : 1; nm++;     // :1381
                    if (P.checkOri) {
                        const int bin = rot_bin(P.kpq[q].angle, P.kpc[bestIdx2].angle);      // :1386-1394
                        P.binOf[q] = bin; P.idxOf[q] = bestIdx2; P.hist[bin]++;
                    }
                }
                __syncwarp();
            }
        } else {
            const int bestDist = best == MD_EMPTY ? 256 : (int)(best >> MD_POSBITS);
            if (bestDist <= MD_TH_HIGH) {                                        // :111
                const int bestIdx = P.cand[p0 + (int)(best & MD_POSMASK)];
                bool ok = true;
                if (second != MD_EMPTY && (second >> MD_POSBITS) < 256u) {        // (a distance of 256 never enters bestDist2, :103)
                    const int bestDist2 = (int)(second >> MD_POSBITS);
                    const int idx2 = P.cand[p0 + (int)(second & MD_POSMASK)];
                    // :113-114  (bestLevel2 stays -1 without a second candidate, and no keypoint has octave -1)
                    if (P.kpc[bestIdx].octave == P.kpc[idx2].octave && (float)bestDist > __fmul_rn(P.nnratio, (float)bestDist2)) ok = false;
                }
                if (ok) {
                    if (lane == 0) { P.out[bestIdx] = q; taken[bestIdx] = 1; nm++; }        // :116-117
                    __syncwarp();
                }
            }
        }
    }
    __syncwarp();
    if (MODE != 0)
        for (int i = lane; i < P.nc; i += 32) P.taken[i] = taken[i];
    if (lane == 0) *P.nmatches = nm;
}

// ComputeThreeMaxima :1554-1595 on the bin counts
__device__ __forceinline__ void three_maxima(const int* hist, int& ind1, int& ind2, int& ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int i = 0; i < MD_HISTO; i++) {
        const int s = hist[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
    else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
}

// SearchForInitialization :481-510: rotation filter, then vbPrevMatched
__global__ void finish_init_kernel(int n1, int checkOri, const int* hist, const int* binOf, int* matches12, int* nmatches, const os1_kp_dev* kp2,
                                   float* prev, const int* status)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n1 || *status) return;
    if (checkOri) {
        int i1, i2, i3;
        three_maxima(hist, i1, i2, i3);
        const int b = binOf[i];
        if (b >= 0 && b != i1 && b != i2 && b != i3 && matches12[i] >= 0) { matches12[i] = -1; atomicSub(nmatches, 1); }
    }
    const int m = matches12[i];
    if (m >= 0) { prev[2 * i] = kp2[m].x; prev[2 * i + 1] = kp2[m].y; }
}

// SearchByProjection(Frame, Frame) :1397-1419: every rotHist entry of a losing bin detaches its keypoint again and counts one match
// less.  (An entry per accepting query, not per keypoint: a keypoint that received a point without observations can be attached
// a second time, and then either entry takes it away, as in the reference.)
__global__ void finish_proj_kernel(int nq, int checkOri, const int* hist, const int* binOf, const int* idxOf, int* match_of_cur, uint8_t* taken,
                                   int* nmatches, const int* status)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq || *status || !checkOri) return;
    int i1, i2, i3;
    three_maxima(hist, i1, i2, i3);
    const int b = binOf[q];
    if (b >= 0 && b != i1 && b != i2 && b != i3) { const int i = idxOf[q]; match_of_cur[i] = OS1_MATCH_REMOVED; taken[i] = 0; atomicSub(nmatches, 1); }
}

size_t claim_smem(int mode, int nc) { return mode == 0 ? (size_t)nc * 4 + 16 : (size_t)nc + 16; }

template <int MODE>
cudaError_t launch_claim(const ClaimParams& P, cudaStream_t s)
{
    const size_t smem = claim_smem(MODE, P.nc);
    if (smem > 48 * 1024) { cudaError_t e = os1_optin_smem((const void*)claim_kernel<MODE>); if (e != cudaSuccess) return e; }
    claim_kernel<MODE><<<1, 32, smem, s>>>(P);
    return cudaGetLastError();
}

GridGeom make_grid(const float* b4)
{
    GridGeom G;
    G.minX = b4[0]; G.minY = b4[2];
    G.invW = (float)OS1_GRID_COLS / (b4[1] - b4[0]);          // mfGridElementWidthInv   Frame.cc:99-100
    G.invH = (float)OS1_GRID_ROWS / (b4[3] - b4[2]);
    return G;
}

int check_sizes(os1_matcher* m, int nq, int nc, const float* b4, const char* who)
{
    if (!m) return mfail(nullptr, OS1_E_INVALID, std::string(who) + ": NULL handle");
    if (nq < 0 || nc < 0 || !b4) return mfail(m, OS1_E_INVALID, std::string(who) + ": bad argument");
    if (nq > m->maxQ || nc > m->maxKp) return mfail(m, OS1_E_CAPACITY, std::string(who) + ": more keypoints / queries than the handle was created for");
    if (!(b4[1] > b4[0]) || !(b4[3] > b4[2])) return mfail(m, OS1_E_INVALID, std::string(who) + ": empty image bounds");
    return OS1_OK;
}

// steps 1, 3, 4 for queries already written to m->q*
int candidates_and_distances(os1_matcher* m, const os1_kp_dev* kpc, int nc, const float* b4, int nq, const uint8_t* Dq, const uint8_t* Dc, cudaStream_t s)
{
    const GridGeom G = make_grid(b4);
    MCK(os1_launch_grid_build(kpc, nc, G, m->cellStart, m->gridIdx, s), "grid build");
    MCK(os1_launch_area_csr(kpc, m->cellStart, m->gridIdx, G, m->qx, m->qy, m->qr, m->qmin, m->qmax, m->qvalid, nq, m->cnt, m->ptr, m->cand, m->maxCand, s),
        "candidate lists");
    // (lists that overflow the capacity are not written; the claim kernel sees ptr[nq] > capacity and raises the status word)
    MCK(os1_launch_hamming_csr(Dq, nq, Dc, m->ptr, m->cand, m->dist, s, m->maxCand), "candidate distances");
    return OS1_OK;
}

} // namespace

extern "C" {

int os1_matcher_create(os1_matcher** out, int max_keypoints, int max_queries, int max_candidates, int device)
{
    if (!out) return mfail(nullptr, OS1_E_INVALID, "os1_matcher_create: out is NULL");
    *out = nullptr;
    if (max_keypoints < 1 || max_queries < 1 || max_candidates < 1) return mfail(nullptr, OS1_E_INVALID, "os1_matcher_create: capacities must be positive");
    if (max_keypoints > OS1_GRID_MAX_KP) return mfail(nullptr, OS1_E_UNSUPPORTED, "os1_matcher_create: at most 16384 keypoints per frame");
    if (max_queries > 32767) return mfail(nullptr, OS1_E_UNSUPPORTED, "os1_matcher_create: at most 32767 queries per call");
    if (max_candidates >= (1 << 30)) return mfail(nullptr, OS1_E_UNSUPPORTED, "os1_matcher_create: candidate capacity too large");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) { cudaGetLastError(); return mfail(nullptr, OS1_E_CUDA, "os1_matcher_create: no CUDA device (there is no CPU fallback)"); }
    if (device < 0 || device >= ndev) return mfail(nullptr, OS1_E_INVALID, "os1_matcher_create: bad device ordinal");
    os1_matcher* m = new os1_matcher();
    m->device = device; m->maxKp = max_keypoints; m->maxQ = max_queries; m->maxCand = max_candidates;
    DeviceScope _ds(device);
    e = _ds.err;
    auto alloc = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes); };
    alloc((void**)&m->cellStart, (OS1_GRID_CELLS + 1) * 4);
    alloc((void**)&m->gridIdx, (size_t)max_keypoints * 4);
    alloc((void**)&m->cnt, (size_t)max_queries * 4);
    alloc((void**)&m->ptr, (size_t)(max_queries + 1) * 4);
    alloc((void**)&m->cand, (size_t)max_candidates * 4);
    alloc((void**)&m->dist, (size_t)max_candidates * 2);
    alloc((void**)&m->qx, (size_t)max_queries * 4); alloc((void**)&m->qy, (size_t)max_queries * 4); alloc((void**)&m->qr, (size_t)max_queries * 4);
    alloc((void**)&m->qmin, (size_t)max_queries * 4); alloc((void**)&m->qmax, (size_t)max_queries * 4);
    alloc((void**)&m->qvalid, (size_t)max_queries);
    alloc((void**)&m->binOf, (size_t)std::max(max_queries, max_keypoints) * 4);
    alloc((void**)&m->idxOf, (size_t)max_queries * 4);
    alloc((void**)&m->hist, 64 * 4);
    alloc((void**)&m->status, 64);
    if (e == cudaSuccess) e = cudaMemset(m->status, 0, 64);
    if (e == cudaSuccess) e = cudaMallocHost((void**)&m->hostStatus, 64);
    if (e == cudaSuccess) e = cudaStreamSynchronize(0);
    if (e != cudaSuccess) {
        cudaGetLastError();
        os1_matcher_destroy(m);
        return mfail(nullptr, OS1_E_CUDA, std::string("os1_matcher_create: ") + cudaGetErrorString(e));
    }
    *out = m;
    return OS1_OK;
}

int os1_matcher_destroy(os1_matcher* m)
{
    if (!m) return OS1_OK;
    DeviceScope _ds(m->device);
    cudaDeviceSynchronize();
    void* bufs[] = { m->cellStart, m->gridIdx, m->cnt, m->ptr, m->cand, m->dist, m->qx, m->qy, m->qr, m->qmin, m->qmax, m->qvalid, m->binOf, m->idxOf, m->hist,
                     m->status, m->stage };
    for (void* p : bufs) if (p) cudaFree(p);
    if (m->hostStatus) cudaFreeHost(m->hostStatus);
    delete m;
    return OS1_OK;
}

const char* os1_matcher_last_error(const os1_matcher* m) { return m ? m->err.c_str() : ""; }

int os1_matcher_synchronize(os1_matcher* m, void* cuda_stream)
{
    if (!m) return mfail(nullptr, OS1_E_INVALID, "os1_matcher_synchronize: NULL handle");
    DeviceScope _ds(m->device); MCK(_ds.err, "cudaSetDevice");
    cudaStream_t s = (cudaStream_t)cuda_stream;
    MCK(cudaMemcpyAsync(m->hostStatus, m->status, 4, cudaMemcpyDeviceToHost, s), "read status");
    MCK(cudaStreamSynchronize(s), "cudaStreamSynchronize");
    if (m->hostStatus[0]) {
        cudaMemsetAsync(m->status, 0, 4, s);
        return mfail(m, OS1_E_CAPACITY, "os1_matcher: the candidate lists of the last call exceeded max_candidates; nothing was matched");
    }
    return OS1_OK;
}

int os1_matcher_search_for_initialization(os1_matcher* m, const os1_keypoint* kp1, const uint8_t* D1, int n1, const os1_keypoint* kp2,
                                          const uint8_t* D2, int n2, const float* bounds4, float* prev_xy, int window_size, float nnratio,
                                          int check_orientation, int32_t* matches12, int32_t* nmatches, void* cuda_stream)
{
    int rc = check_sizes(m, n1, n2, bounds4, "os1_matcher_search_for_initialization");
    if (rc) return rc;
    if (!nmatches || (n1 > 0 && (!kp1 || !D1 || !prev_xy || !matches12)) || (n2 > 0 && (!kp2 || !D2)))
        return mfail(m, OS1_E_INVALID, "os1_matcher_search_for_initialization: NULL argument");
    DeviceScope _ds(m->device); MCK(_ds.err, "cudaSetDevice");
    cudaStream_t s = (cudaStream_t)cuda_stream;
    if (n1 == 0) { MCK(cudaMemsetAsync(nmatches, 0, 4, s), "cudaMemsetAsync"); return OS1_OK; }
    init_query_kernel<<<(n1 + 127) / 128, 128, 0, s>>>((const os1_kp_dev*)kp1, n1, prev_xy, (float)window_size, m->qx, m->qy, m->qr, m->qmin, m->qmax, m->qvalid);
    rc = candidates_and_distances(m, (const os1_kp_dev*)kp2, n2, bounds4, n1, D1, D2, s);
    if (rc) return rc;
    ClaimParams P;
    memset(&P, 0, sizeof(P));
    P.nq = n1; P.nc = n2; P.ptr = m->ptr; P.cand = m->cand; P.dist = m->dist; P.candCap = m->maxCand; P.nnratio = nnratio; P.checkOri = check_orientation != 0;
    P.kpq = (const os1_kp_dev*)kp1; P.kpc = (const os1_kp_dev*)kp2; P.out = matches12; P.nmatches = nmatches; P.binOf = m->binOf; P.hist = m->hist;
    P.status = m->status;
    MCK(launch_claim<0>(P, s), "claim kernel");
    finish_init_kernel<<<(n1 + 127) / 128, 128, 0, s>>>(n1, P.checkOri, m->hist, m->binOf, matches12, nmatches, (const os1_kp_dev*)kp2, prev_xy, m->status);
    MCK(cudaGetLastError(), "finish kernel");
    return OS1_OK;
}

static int fill_proj(os1_matcher* m, ProjParams& P, const float* Tcw16, const float* K4, const float* b4, const float* scale, int nlevels, float th,
                     const char* who)
{
    if (nlevels < 1 || nlevels > OS1_MAX_LEVELS || !scale) return mfail(m, OS1_E_INVALID, std::string(who) + ": bad scale-factor table");
    memset(&P, 0, sizeof(P));
    if (Tcw16) memcpy(P.T, Tcw16, 12 * sizeof(float));
    if (K4) { P.fx = K4[0]; P.fy = K4[1]; P.cx = K4[2]; P.cy = K4[3]; }
    P.minX = b4[0]; P.maxX = b4[1]; P.minY = b4[2]; P.maxY = b4[3];
    P.th = th; P.nlevels = nlevels;
    for (int i = 0; i < nlevels; i++) P.scale[i] = scale[i];
    return OS1_OK;
}

int os1_matcher_search_by_projection_frame(os1_matcher* m, const os1_keypoint* kp_cur, const uint8_t* D_cur, int n_cur, const float* Tcw_cur,
                                           uint8_t* taken, const os1_keypoint* kp_last, int n_last, const uint8_t* has_point,
                                           const uint8_t* outlier, const float* world_pos, const uint8_t* D_points, const float* K4,
                                           const float* bounds4, const float* scale_factors, int nlevels, float th, int check_orientation,
                                           int32_t* match_of_cur, int32_t* nmatches, void* cuda_stream)
{
    int rc = check_sizes(m, n_last, n_cur, bounds4, "os1_matcher_search_by_projection_frame");
    if (rc) return rc;
    if (!nmatches || !Tcw_cur || !K4 || (n_cur > 0 && (!kp_cur || !D_cur || !taken || !match_of_cur)) ||
        (n_last > 0 && (!kp_last || !has_point || !outlier || !world_pos || !D_points)))
        return mfail(m, OS1_E_INVALID, "os1_matcher_search_by_projection_frame: NULL argument");
    ProjParams PP;
    rc = fill_proj(m, PP, Tcw_cur, K4, bounds4, scale_factors, nlevels, th, "os1_matcher_search_by_projection_frame");
    if (rc) return rc;
    DeviceScope _ds(m->device); MCK(_ds.err, "cudaSetDevice");
    cudaStream_t s = (cudaStream_t)cuda_stream;
    if (n_cur == 0) { MCK(cudaMemsetAsync(nmatches, 0, 4, s), "cudaMemsetAsync"); return OS1_OK; }
    if (n_last > 0)
        project_query_kernel<<<(n_last + 127) / 128, 128, 0, s>>>((const os1_kp_dev*)kp_last, n_last, has_point, outlier, world_pos, PP, m->qx, m->qy, m->qr,
                                                                 m->qmin, m->qmax, m->qvalid);
    rc = candidates_and_distances(m, (const os1_kp_dev*)kp_cur, n_cur, bounds4, n_last, D_points, D_cur, s);
    if (rc) return rc;
    ClaimParams P;
    memset(&P, 0, sizeof(P));
    P.nq = n_last; P.nc = n_cur; P.ptr = m->ptr; P.cand = m->cand; P.dist = m->dist; P.candCap = m->maxCand; P.checkOri = check_orientation != 0;
    P.kpq = (const os1_kp_dev*)kp_last; P.kpc = (const os1_kp_dev*)kp_cur; P.taken = taken; P.qhas = has_point; P.out = match_of_cur; P.nmatches = nmatches;
    P.binOf = m->binOf; P.idxOf = m->idxOf; P.hist = m->hist; P.status = m->status;
    MCK(launch_claim<1>(P, s), "claim kernel");
    if (n_last > 0)
        finish_proj_kernel<<<(n_last + 127) / 128, 128, 0, s>>>(n_last, P.checkOri, m->hist, m->binOf, m->idxOf, match_of_cur, taken, nmatches, m->status);
    MCK(cudaGetLastError(), "finish kernel");
    return OS1_OK;
}

int os1_matcher_search_by_projection_points(os1_matcher* m, const os1_keypoint* kp_frame, const uint8_t* D_frame, int n_frame, uint8_t* taken,
                                            const float* map_points, const uint8_t* D_points, int n_points, const float* bounds4,
                                            const float* scale_factors, int nlevels, float th, float nnratio, int32_t* match_of_frame,
                                            int32_t* nmatches, void* cuda_stream)
{
    int rc = check_sizes(m, n_points, n_frame, bounds4, "os1_matcher_search_by_projection_points");
    if (rc) return rc;
    if (!nmatches || (n_frame > 0 && (!kp_frame || !D_frame || !taken || !match_of_frame)) || (n_points > 0 && (!map_points || !D_points)))
        return mfail(m, OS1_E_INVALID, "os1_matcher_search_by_projection_points: NULL argument");
    ProjParams PP;
    rc = fill_proj(m, PP, nullptr, nullptr, bounds4, scale_factors, nlevels, th, "os1_matcher_search_by_projection_points");
    if (rc) return rc;
    DeviceScope _ds(m->device); MCK(_ds.err, "cudaSetDevice");
    cudaStream_t s = (cudaStream_t)cuda_stream;
    if (n_frame == 0) { MCK(cudaMemsetAsync(nmatches, 0, 4, s), "cudaMemsetAsync"); return OS1_OK; }
    if (n_points > 0)
        point_query_kernel<<<(n_points + 127) / 128, 128, 0, s>>>(map_points, n_points, PP, m->qx, m->qy, m->qr, m->qmin, m->qmax, m->qvalid);
    rc = candidates_and_distances(m, (const os1_kp_dev*)kp_frame, n_frame, bounds4, n_points, D_points, D_frame, s);
    if (rc) return rc;
    ClaimParams P;
    memset(&P, 0, sizeof(P));
    P.nq = n_points; P.nc = n_frame; P.ptr = m->ptr; P.cand = m->cand; P.dist = m->dist; P.candCap = m->maxCand; P.nnratio = nnratio;
    P.kpc = (const os1_kp_dev*)kp_frame; P.taken = taken; P.out = match_of_frame; P.nmatches = nmatches; P.binOf = m->binOf; P.hist = m->hist;
    P.status = m->status;
    MCK(launch_claim<2>(P, s), "claim kernel");
    return OS1_OK;
}

} // extern "C"

// ---- the same three calls with HOST buffers (what the C++ adapter include/ORBmatcher.h uses): inputs are staged into a
// device buffer the handle keeps (grown when a call needs more), the device pipeline runs, results come back, the call waits.
namespace {

struct Stager {
    os1_matcher* m; cudaStream_t s; size_t off = 0; cudaError_t e = cudaSuccess;
    size_t need = 0;
    static size_t al(size_t b) { return (b + 255) & ~(size_t)255; }
    void want(size_t bytes) { need += al(std::max<size_t>(bytes, 1)); }
    cudaError_t reserve()
    {
        if (need <= m->stageBytes) return cudaSuccess;
        cudaError_t r = cudaStreamSynchronize(s);
        if (r != cudaSuccess) return r;
        if (m->stage) cudaFree(m->stage);
        m->stage = nullptr; m->stageBytes = 0;
        r = cudaMalloc((void**)&m->stage, need);
        if (r == cudaSuccess) m->stageBytes = need;
        return r;
    }
    template <typename T> T* put(const T* host, size_t count)
    {
        T* d = (T*)(m->stage + off);
        off += al(std::max<size_t>(count * sizeof(T), 1));
        if (e == cudaSuccess && host && count) e = cudaMemcpyAsync(d, host, count * sizeof(T), cudaMemcpyHostToDevice, s);
        return d;
    }
    template <typename T> void get(T* host, const T* dev, size_t count)
    {
        if (e == cudaSuccess && host && count) e = cudaMemcpyAsync(host, dev, count * sizeof(T), cudaMemcpyDeviceToHost, s);
    }
};

} // namespace

extern "C" {

int os1_matcher_search_for_initialization_host(os1_matcher* m, const os1_keypoint* kp1, const uint8_t* D1, int n1, const os1_keypoint* kp2,
                                               const uint8_t* D2, int n2, const float* bounds4, float* prev_xy, int window_size, float nnratio,
                                               int check_orientation, int32_t* matches12, int32_t* nmatches)
{
    int rc = check_sizes(m, n1, n2, bounds4, "os1_matcher_search_for_initialization_host");
    if (rc) return rc;
    if (!nmatches || (n1 > 0 && (!kp1 || !D1 || !prev_xy || !matches12)) || (n2 > 0 && (!kp2 || !D2)))
        return mfail(m, OS1_E_INVALID, "os1_matcher_search_for_initialization_host: NULL argument");
    DeviceScope _ds(m->device); MCK(_ds.err, "cudaSetDevice");
    Stager st{ m, nullptr };
    st.want((size_t)n1 * sizeof(os1_keypoint)); st.want((size_t)n1 * 32); st.want((size_t)n2 * sizeof(os1_keypoint)); st.want((size_t)n2 * 32);
    st.want((size_t)n1 * 8); st.want((size_t)n1 * 4); st.want(4);
    MCK(st.reserve(), "staging buffer");
    const os1_keypoint* dk1 = st.put(kp1, n1); const uint8_t* dD1 = st.put(D1, (size_t)n1 * 32);
    const os1_keypoint* dk2 = st.put(kp2, n2); const uint8_t* dD2 = st.put(D2, (size_t)n2 * 32);
    float* dprev = st.put(prev_xy, (size_t)n1 * 2);
    int32_t* dm = st.put((const int32_t*)nullptr, n1); int32_t* dn = st.put((const int32_t*)nullptr, 1);
    MCK(st.e, "upload");
    rc = os1_matcher_search_for_initialization(m, dk1, dD1, n1, dk2, dD2, n2, bounds4, dprev, window_size, nnratio, check_orientation, dm, dn, nullptr);
    if (rc) return rc;
    st.get(prev_xy, dprev, (size_t)n1 * 2); st.get(matches12, dm, n1); st.get(nmatches, dn, 1);
    MCK(st.e, "download");
    return os1_matcher_synchronize(m, nullptr);
}

int os1_matcher_search_by_projection_frame_host(os1_matcher* m, const os1_keypoint* kp_cur, const uint8_t* D_cur, int n_cur, const float* Tcw_cur,
                                                uint8_t* taken, const os1_keypoint* kp_last, int n_last, const uint8_t* has_point,
                                                const uint8_t* outlier, const float* world_pos, const uint8_t* D_points, const float* K4,
                                                const float* bounds4, const float* scale_factors, int nlevels, float th, int check_orientation,
                                                int32_t* match_of_cur, int32_t* nmatches)
{
    int rc = check_sizes(m, n_last, n_cur, bounds4, "os1_matcher_search_by_projection_frame_host");
    if (rc) return rc;
    if (!nmatches || (n_cur > 0 && (!kp_cur || !D_cur || !taken || !match_of_cur)) ||
        (n_last > 0 && (!kp_last || !has_point || !outlier || !world_pos || !D_points)))
        return mfail(m, OS1_E_INVALID, "os1_matcher_search_by_projection_frame_host: NULL argument");
    DeviceScope _ds(m->device); MCK(_ds.err, "cudaSetDevice");
    Stager st{ m, nullptr };
    st.want((size_t)n_cur * sizeof(os1_keypoint)); st.want((size_t)n_cur * 32); st.want(n_cur); st.want((size_t)n_last * sizeof(os1_keypoint));
    st.want(n_last); st.want(n_last); st.want((size_t)n_last * 12); st.want((size_t)n_last * 32); st.want((size_t)n_cur * 4); st.want(4);
    MCK(st.reserve(), "staging buffer");
    const os1_keypoint* dkc = st.put(kp_cur, n_cur); const uint8_t* dDc = st.put(D_cur, (size_t)n_cur * 32); uint8_t* dtk = st.put(taken, n_cur);
    const os1_keypoint* dkl = st.put(kp_last, n_last); const uint8_t* dhas = st.put(has_point, n_last); const uint8_t* dout = st.put(outlier, n_last);
    const float* dw = st.put(world_pos, (size_t)n_last * 3); const uint8_t* dDp = st.put(D_points, (size_t)n_last * 32);
    int32_t* dm = st.put((const int32_t*)nullptr, n_cur); int32_t* dn = st.put((const int32_t*)nullptr, 1);
    MCK(st.e, "upload");
    rc = os1_matcher_search_by_projection_frame(m, dkc, dDc, n_cur, Tcw_cur, dtk, dkl, n_last, dhas, dout, dw, dDp, K4, bounds4, scale_factors, nlevels, th,
                                                check_orientation, dm, dn, nullptr);
    if (rc) return rc;
    st.get(taken, dtk, n_cur); st.get(match_of_cur, dm, n_cur); st.get(nmatches, dn, 1);
    MCK(st.e, "download");
    return os1_matcher_synchronize(m, nullptr);
}

int os1_matcher_search_by_projection_points_host(os1_matcher* m, const os1_keypoint* kp_frame, const uint8_t* D_frame, int n_frame, uint8_t* taken,
                                                 const float* map_points, const uint8_t* D_points, int n_points, const float* bounds4,
                                                 const float* scale_factors, int nlevels, float th, float nnratio, int32_t* match_of_frame,
                                                 int32_t* nmatches)
{
    int rc = check_sizes(m, n_points, n_frame, bounds4, "os1_matcher_search_by_projection_points_host");
    if (rc) return rc;
    if (!nmatches || (n_frame > 0 && (!kp_frame || !D_frame || !taken || !match_of_frame)) || (n_points > 0 && (!map_points || !D_points)))
        return mfail(m, OS1_E_INVALID, "os1_matcher_search_by_projection_points_host: NULL argument");
    DeviceScope _ds(m->device); MCK(_ds.err, "cudaSetDevice");
    Stager st{ m, nullptr };
    st.want((size_t)n_frame * sizeof(os1_keypoint)); st.want((size_t)n_frame * 32); st.want(n_frame); st.want((size_t)n_points * 24);
    st.want((size_t)n_points * 32); st.want((size_t)n_frame * 4); st.want(4);
    MCK(st.reserve(), "staging buffer");
    const os1_keypoint* dk = st.put(kp_frame, n_frame); const uint8_t* dD = st.put(D_frame, (size_t)n_frame * 32); uint8_t* dtk = st.put(taken, n_frame);
    const float* dmp = st.put(map_points, (size_t)n_points * 6); const uint8_t* dDp = st.put(D_points, (size_t)n_points * 32);
    int32_t* dm = st.put((const int32_t*)nullptr, n_frame); int32_t* dn = st.put((const int32_t*)nullptr, 1);
    MCK(st.e, "upload");
    rc = os1_matcher_search_by_projection_points(m, dk, dD, n_frame, dtk, dmp, dDp, n_points, bounds4, scale_factors, nlevels, th, nnratio, dm, dn, nullptr);
    if (rc) return rc;
    st.get(taken, dtk, n_frame); st.get(match_of_frame, dm, n_frame); st.get(nmatches, dn, 1);
    MCK(st.e, "download");
    return os1_matcher_synchronize(m, nullptr);
}

} // extern "C"
