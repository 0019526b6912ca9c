// C ABI of os1_b200 (include/os1_b200.h): handle management, geometry/tables that the reference computes in
// its constructor and at the top of each stage, buffer ownership, and the four-kernel extraction pipeline.
#include "../../include/os1_b200.h"
#include "os1_internal.cuh"
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace {

thread_local std::string t_err;

struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    cudaError_t ensure(size_t n)
    {
        if (n <= bytes) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; bytes = 0;
        cudaError_t e = cudaMalloc(&p, n);
        // slots a frame does not fill read as zero, never as stale data.  The memset runs on the legacy stream, which the
        // handle's non-blocking streams do not wait for: finish it here (set-up path) so that it can never land on top
        // of an upload or a result queued right after
        if (e == cudaSuccess) { bytes = n; e = cudaMemset(p, 0, n); }
        if (e == cudaSuccess) e = cudaStreamSynchronize(0);
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
};

} // namespace

struct os1_orb {
    // constructor state (ORBextractor.cc:442-502)
    int nfeatures, nlevels, iniTh, minTh;
    double scaleFactor;      // ORBextractor.h:313: a double that holds (double)(float)argument
    std::vector<float> scale, inv, sigma2, invsigma2;
    std::vector<int> quota;
    int maxQuota;

    int device;
    cudaStream_t ownStream = nullptr, stream = nullptr;
    cudaStream_t inStream = nullptr, outStream = nullptr;       // upload / download streams of the chunked host path
    cudaStream_t altStream[3] = { nullptr, nullptr, nullptr };  // extra compute streams: chunks rotate over them, so one chunk's
    cudaEvent_t altEv[3] = { nullptr, nullptr, nullptr };       // partial last waves overlap the next chunk's kernels
    std::vector<cudaEvent_t> chunkEv;                           // 2 per chunk: uploaded, computed
    std::string err;

    // geometry of the current resolution
    bool haveGeom = false;
    Geometry geom;
    int maxCells = 0;
    int batchCap = 0;
    // latency path: the four launches and the result copy of a single host frame as one CUDA graph (built on the third
    // call of a geometry, dropped whenever a buffer it points to is reallocated)
    cudaGraphExec_t latExec = nullptr;
    int latCalls = 0;
    bool latDisabled = false;
    PyramidPlan* plan = nullptr;        // small tiles (single frames)
    PyramidPlan* planBig = nullptr;     // larger tiles (batches)
    DevBuf rtab, ptab, gtab, slab, cellCnt, stage, cand, code, node, candCount, lvKp, lvCount, status, outKp, outDesc, outCount;
    int lastFrames = 0;
    int lastLaunches = 4;       // kernels the last extract call launched
    const uint8_t* lastLvl0 = nullptr; long long lastLvl0Stride = 0; int lastLvl0Pitch = 0;
    // pinned host mirrors for the host-buffer path
    void* hostOut = nullptr; size_t hostOutBytes = 0;
    int* hostStatus = nullptr;
    // optional per-kernel timing (bench.py's roofline): events around the four launches of the last call
    bool timing = false;
    cudaEvent_t ev[5] = { nullptr, nullptr, nullptr, nullptr, nullptr };
};

namespace {

int fail(os1_orb* h, int code, const std::string& msg)
{
    if (h) h->err = msg;
    t_err = msg;
    return code;
}
int cuda_fail(os1_orb* h, cudaError_t e, const char* what)
{
    cudaGetLastError();
    return fail(h, OS1_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CKD(call, what)                                                       \
    do {                                                                      \
        cudaError_t _e = (call);                                              \
        if (_e != cudaSuccess) { drain(); return cuda_fail(h, _e, what); }    \
    } while (0)
#define CK(call, what)                                          \
    do {                                                        \
        cudaError_t _e = (call);                                \
        if (_e != cudaSuccess) return cuda_fail(h, _e, what);   \
    } while (0)

inline int cv_round(float v) { return (int)lrintf(v); }
inline int cv_floor(float v) { int i = (int)v; return i - (i > v); }

// scale tables and per-level quotas exactly as the reference constructor computes them (:447-478)
void build_tables(os1_orb* h)
{
    const int L = h->nlevels;
    h->scale.assign(L, 0); h->sigma2.assign(L, 0); h->inv.assign(L, 0); h->invsigma2.assign(L, 0); h->quota.assign(L, 0);
    h->scale[0] = 1.0f; h->sigma2[0] = 1.0f;
    for (int i = 1; i < L; i++) {
        h->scale[i] = (float)(h->scale[i - 1] * h->scaleFactor);
        h->sigma2[i] = h->scale[i] * h->scale[i];
    }
    for (int i = 0; i < L; i++) { h->inv[i] = 1.0f / h->scale[i]; h->invsigma2[i] = 1.0f / h->sigma2[i]; }
    float factor = (float)(1.0f / h->scaleFactor);
    float per = h->nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)L));
    int sum = 0;
    for (int l = 0; l < L - 1; l++) {
        h->quota[l] = cv_round(per);
        sum += h->quota[l];
        per *= factor;
    }
    h->quota[L - 1] = std::max(h->nfeatures - sum, 0);
    h->maxQuota = 0;
    for (int l = 0; l < L; l++) h->maxQuota = std::max(h->maxQuota, h->quota[l]);
}

short sat_short(float v)
{
    int i = cv_round(v);
    return (short)(i < -32768 ? -32768 : i > 32767 ? 32767 : i);
}

} // namespace

// lets the other translation units of the library report through os1_last_error(NULL)
void os1_internal_set_error(const char* msg) { t_err = msg ? msg : ""; }

cudaError_t os1_optin_smem(const void* kernel)
{
    static std::mutex mu;
    static std::vector<std::pair<const void*, int> > done;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lk(mu);
    for (const auto& d : done)
        if (d.first == kernel && d.second == dev) return cudaSuccess;
    int optin = 0;
    cudaFuncAttributes fa;
    e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&fa, kernel);
    // (the opt-in limit covers static + dynamic shared memory together)
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - (int)fa.sharedSizeBytes);
    if (e == cudaSuccess) done.push_back(std::make_pair(kernel, dev));
    return e;
}

// cv::resize INTER_LINEAR taps of one axis: first/second source index (already clamped
// the way cv::resize clamps them on that axis) and the two 11-bit coefficients.
void os1_host_resize_taps(int ssize, int dsize, bool vertical, std::vector<short>& s0, std::vector<short>& s1, std::vector<short>& c0,
                          std::vector<short>& c1)
{
    const double inv_scale = (double)dsize / ssize;
    const double scale = 1. / inv_scale;
    s0.resize(dsize); s1.resize(dsize); c0.resize(dsize); c1.resize(dsize);
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cv_floor(f);
        f -= s;
        if (!vertical) {
            // horizontal: out-of-range taps reset the weight (xofs / alpha tables)
            if (s < 0) { f = 0; s = 0; }
            if (s >= ssize - 1) { f = 0; s = ssize - 1; }
            s0[d] = (short)s; s1[d] = (short)std::min(s + 1, ssize - 1);
        } else {
            // vertical: the weight is kept and the two ROW INDICES are clamped
            s0[d] = (short)std::min(std::max(s, 0), ssize - 1);
            s1[d] = (short)std::min(std::max(s + 1, 0), ssize - 1);
        }
        c0[d] = sat_short((1.f - f) * 2048.f);
        c1[d] = sat_short(f * 2048.f);
    }
}

namespace {

// DivideNode's split bits along one axis for every coordinate (ORBextractor.cc:515-516, :547-557), spread to
// every second bit: depth d (1-based) sits at bit 2*(PATH_BITS-d) (+1 for y).  For x the root index
// (:584-601) goes above the path bits.
void build_path_table(int extent, int nIni, bool isX, std::vector<uint32_t>& out)
{
    const int D = OS1_PATH_BITS;
    out.resize(extent);
    const float hX = (float)extent / nIni;
    for (int p = 0; p < extent; p++) {
        int a = 0, b = extent;
        uint32_t code = 0;
        if (isX) {
            int r = (int)((float)p / hX);
            if (r >= nIni) r = nIni - 1;
            a = (int)(hX * (float)r);
            b = (int)(hX * (float)(r + 1));
            code = (uint32_t)r << (2 * D);
        }
        for (int d = 1; d <= D; d++) {
            const int half = (int)ceil((float)(b - a) / 2);
            const int mid = a + half;
            const int bit = ((float)p < (float)mid) ? 0 : 1;
            if (bit) a = mid; else b = mid;
            code |= (uint32_t)bit << (2 * (D - d) + (isX ? 0 : 1));
        }
        out[p] = code;
    }
}

int build_geometry(os1_orb* h, int rows, int cols)
{
    Geometry& g = h->geom;
    memset(&g, 0, sizeof(g));
    const int L = h->nlevels;
    g.nlevels = L; g.rows = rows; g.cols = cols; g.iniTh = h->iniTh; g.minTh = h->minTh;
    long long slab = 0, candOff = 0;
    int cellOff = 0, groupOff = 0, kpOff = 0;
    std::vector<short4> rtab;
    std::vector<uint32_t> ptab;
    h->maxCells = 0;
    for (int l = 0; l < L; l++) {
        LevelGeom& v = g.lv[l];
        v.w = cv_round((float)cols * h->inv[l]);                          // :976
        v.h = cv_round((float)rows * h->inv[l]);
        if (v.w > 4096 + 32 || v.h > 4096 + 32) return fail(h, OS1_E_UNSUPPORTED, "image side above 4128 px");
        v.pitch = (v.w + 63) & ~63;
        v.off = slab;
        slab += (long long)v.pitch * v.h;
        slab = (slab + 255) & ~255LL;
        v.maxBX = v.w - OS1_EDGE + 3; v.maxBY = v.h - OS1_EDGE + 3;     // :809-810
        v.W = v.maxBX - OS1_BORDER; v.H = v.maxBY - OS1_BORDER;
        const float width = (float)v.W, height = (float)v.H;
        v.nCols = (int)(width / 30.f); v.nRows = (int)(height / 30.f);   // :820-821
        if (v.w < 1 || v.h < 1) return fail(h, OS1_E_UNSUPPORTED, "a pyramid level has zero size (cv::resize throws there)");
        // A level narrower or lower than 62 px has no FAST cell row/column: the reference's cell loops do not run
        // (:826,:837) and DistributeOctTree gets nothing, so the level is built but yields no keypoints.
        const bool dead = v.nCols < 1 || v.nRows < 1;
        if (dead) { v.nCols = v.nRows = 0; v.wCell = v.hCell = 1; v.W = std::max(v.W, 1); v.H = std::max(v.H, 1); }
        else { v.wCell = (int)ceil(width / v.nCols); v.hCell = (int)ceil(height / v.nRows); }   // :822-823
        v.nCells = v.nCols * v.nRows;
        v.cellOff = cellOff; cellOff += v.nCells;
        v.cellCap = ((v.wCell + 1) / 2) * ((v.hCell + 1) / 2);
        v.candCap = v.nCells * v.cellCap;
        if (v.candCap >= (1 << 24)) return fail(h, OS1_E_UNSUPPORTED, "more than 2^24 candidate slots in one level");
        v.candOff = candOff; candOff += v.candCap;
        v.groupsPerRow = (v.nCols + FAST_G - 1) / FAST_G;
        v.groupOff = groupOff; groupOff += v.groupsPerRow * v.nRows;
        h->maxCells = std::max(h->maxCells, v.nCells);
        v.nIni = dead ? 1 : (int)roundf((float)v.W / v.H);                // :575
        if (v.nIni < 1) return fail(h, OS1_E_UNSUPPORTED, "portrait images taller than 2:1 (the reference's nIni is 0 there)");
        if (v.nIni > 4) return fail(h, OS1_E_UNSUPPORTED, "images wider than 4.5:1");
        v.quota = h->quota[l];
        v.xsOff = (int)ptab.size();
        { std::vector<uint32_t> t; build_path_table(v.W, v.nIni, true, t); ptab.insert(ptab.end(), t.begin(), t.end()); }
        v.ysOff = (int)ptab.size();
        { std::vector<uint32_t> t; build_path_table(v.H, 1, false, t); ptab.insert(ptab.end(), t.begin(), t.end()); }
        v.kpCap = std::max(v.quota + 2, 4 * v.nIni) + 2;
        v.kpOff = kpOff; kpOff += v.kpCap;
        v.scale = h->scale[l];
        v.kpSize = (float)(int)(31 * h->scale[l]);                         // :877
        if (l > 0) {
            std::vector<short> s0, s1, c0, c1;
            v.rxOff = (int)rtab.size();
            os1_host_resize_taps(g.lv[l - 1].w, v.w, false, s0, s1, c0, c1);
            for (int i = 0; i < v.w; i++) rtab.push_back(make_short4(s0[i], s1[i], c0[i], c1[i]));
            v.ryOff = (int)rtab.size();
            os1_host_resize_taps(g.lv[l - 1].h, v.h, true, s0, s1, c0, c1);
            for (int i = 0; i < v.h; i++) rtab.push_back(make_short4(s0[i], s1[i], c0[i], c1[i]));
            if (g.lv[l - 1].w == 2 * v.w && g.lv[l - 1].h == 2 * v.h)
                return fail(h, OS1_E_UNSUPPORTED, "scaleFactor 2.0: cv::resize silently switches to INTER_AREA there (not implemented)");
        }
    }
    {
        int maxIni = 1;
        for (int l = 0; l < L; l++) maxIni = std::max(maxIni, g.lv[l].nIni);
        if (!os1_quadtree_fits(h->maxQuota, h->maxCells, maxIni))
            return fail(h, OS1_E_UNSUPPORTED, "too many features in one pyramid level for this image size: the quadtree state of a level must fit one CTA's shared memory (about 2300 features per level)");
    }
    g.slabBytes = slab;
    g.cellsPerFrame = cellOff;
    g.groupsPerFrame = groupOff;
    g.kpCapPerFrame = kpOff;
    g.candPerFrame = candOff;
    if (rtab.empty()) rtab.push_back(make_short4(0, 0, 0, 0));
    CK(h->rtab.ensure(rtab.size() * sizeof(short4)), "cudaMalloc(rtab)");
    CK(h->ptab.ensure(ptab.size() * sizeof(uint32_t)), "cudaMalloc(ptab)");
    CK(cudaMemcpy(h->rtab.p, rtab.data(), rtab.size() * sizeof(short4), cudaMemcpyHostToDevice), "upload resize tables");
    CK(cudaMemcpy(h->ptab.p, ptab.data(), ptab.size() * sizeof(uint32_t), cudaMemcpyHostToDevice), "upload path tables");
    {
        // one entry per FAST CTA of a frame, so the kernel neither searches its level nor divides by the groups per row
        std::vector<int2> gtab;
        for (int l = 0; l < L; l++)
            for (int ci = 0; ci < g.lv[l].nRows; ci++)
                for (int gj = 0; gj < g.lv[l].groupsPerRow; gj++) gtab.push_back(make_int2(l | (ci << 8), gj * FAST_G));
        if (gtab.empty()) gtab.push_back(make_int2(0, 0));
        CK(h->gtab.ensure(gtab.size() * sizeof(int2)), "cudaMalloc(gtab)");
        CK(cudaMemcpy(h->gtab.p, gtab.data(), gtab.size() * sizeof(int2), cudaMemcpyHostToDevice), "upload FAST group table");
    }
    if (h->plan) os1_pyramid_plan_destroy(h->plan);
    // coarsest-level tile: small for single frames (more CTAs), the plan is per resolution
    if (h->planBig) os1_pyramid_plan_destroy(h->planBig);
    h->planBig = nullptr;
    {
        // coarsest-level tile: 32x16 keeps a single frame spread over enough CTAs.  A tile whose level-0 footprint exceeds a
        // TMA box (256) falls back to a smaller one.
        const int cand[][2] = { { 32, 16 }, { 24, 12 }, { 16, 8 }, { 8, 8 } };
        for (auto& c : cand) {
            h->plan = os1_pyramid_plan_create(g, c[0], c[1]);
            if (os1_pyramid_plan_ok(h->plan)) break;
            os1_pyramid_plan_destroy(h->plan); h->plan = nullptr;
        }
        if (!h->plan) return fail(h, OS1_E_UNSUPPORTED, "no pyramid tile fits a TMA box for this scale factor / level count");
        // batches: the largest tile whose level-0 footprint still fits a TMA box (52x26 at scale 1.2 / 8 levels: 0.62 ms per
        // 256 x 720p against 0.68 for 40x20 — less halo, fewer block-wide phases per pixel)
        int big[][2] = { { 52, 26 }, { 40, 20 }, { 32, 16 }, { 24, 12 } };
        if (const char* e = getenv("OS1_PYR_TILE")) sscanf(e, "%dx%d", &big[0][0], &big[0][1]);
        for (auto& c : big) {
            h->planBig = os1_pyramid_plan_create(g, c[0], c[1]);
            if (os1_pyramid_plan_ok(h->planBig)) break;
            os1_pyramid_plan_destroy(h->planBig); h->planBig = nullptr;
        }
    }
    CK(os1_describe_init(), "describe tables");
    // the table uploads above are pageable-memory copies on the legacy stream; the handle's streams are non-blocking and
    // would not wait for them
    CK(cudaStreamSynchronize(0), "sync after table uploads");
    h->haveGeom = true;
    h->batchCap = 0;
    return OS1_OK;
}

int ensure(os1_orb* h, int rows, int cols, int nframes)
{
    if (!h->haveGeom || h->geom.rows != rows || h->geom.cols != cols) {
        if (h->haveGeom) CK(cudaStreamSynchronize(h->stream), "sync before re-planning");
        if (h->latExec) { cudaGraphExecDestroy(h->latExec); h->latExec = nullptr; }
        h->latCalls = 0;
        int rc = build_geometry(h, rows, cols);
        if (rc) return rc;
    }
    if (nframes > h->batchCap) {
        const Geometry& g = h->geom;
        CK(cudaStreamSynchronize(h->stream), "sync before growing buffers");
        if (h->latExec) { cudaGraphExecDestroy(h->latExec); h->latExec = nullptr; }
        h->latCalls = 0;
        const size_t B = (size_t)nframes;
        CK(h->slab.ensure(B * g.slabBytes + 256), "cudaMalloc(pyramid slab)");
        CK(h->cellCnt.ensure(B * g.cellsPerFrame * 4), "cudaMalloc(cell counts)");
        CK(h->stage.ensure(B * g.candPerFrame * 4), "cudaMalloc(candidate staging)");
        CK(h->cand.ensure(B * g.candPerFrame * 4), "cudaMalloc(candidates)");
        CK(h->code.ensure(B * g.candPerFrame * 4), "cudaMalloc(path codes)");
        CK(h->node.ensure(B * g.candPerFrame * 2 + 16), "cudaMalloc(node labels)");
        CK(h->candCount.ensure(B * g.nlevels * 4), "cudaMalloc(candidate counts)");
        CK(h->lvKp.ensure(B * g.kpCapPerFrame * 4), "cudaMalloc(level keypoints)");
        CK(h->lvCount.ensure(B * g.nlevels * 4), "cudaMalloc(level counts)");
        // status word (offset 0), then the single-frame outputs — count (offset 4), keypoints (offset 64), descriptors — so
        // that the latency path fetches everything with ONE device-to-host copy
        CK(h->status.ensure(64 + (size_t)g.kpCapPerFrame * (sizeof(os1_kp_dev) + 32)), "cudaMalloc(status + single-frame outputs)");
        CK(cudaMemset(h->status.p, 0, 64), "cudaMemset(status)");
        CK(cudaStreamSynchronize(0), "sync after cudaMemset(status)");
        CK(h->outKp.ensure(B * g.kpCapPerFrame * sizeof(os1_kp_dev)), "cudaMalloc(keypoints)");
        CK(h->outDesc.ensure(B * g.kpCapPerFrame * 32), "cudaMalloc(descriptors)");
        CK(h->outCount.ensure(B * 4), "cudaMalloc(counts)");
        const size_t hostBytes = B * g.kpCapPerFrame * (sizeof(os1_kp_dev) + 32) + B * 4 + 64;
        if (hostBytes > h->hostOutBytes) {
            if (h->hostOut) cudaFreeHost(h->hostOut);
            h->hostOut = nullptr; h->hostOutBytes = 0;
            CK(cudaMallocHost(&h->hostOut, hostBytes), "cudaMallocHost(outputs)");
            h->hostOutBytes = hostBytes;
        }
        if (!h->hostStatus) CK(cudaMallocHost((void**)&h->hostStatus, 64), "cudaMallocHost(status)");
        h->batchCap = nframes;
    }
    return OS1_OK;
}

int check_status(os1_orb* h, int st)
{
    if (!st) return OS1_OK;
    std::string m = "device-side capacity error:";
    if (st & OS1_ERR_NODE_OVERFLOW) m += " quadtree node table overflow;";
    if (st & OS1_ERR_DEPTH) m += " quadtree deeper than path code;";
    if (st & OS1_ERR_KP_OVERFLOW) m += " keypoint capacity exceeded;";
    if (st & OS1_ERR_CELL_OVERFLOW) m += " FAST cell staging overflow;";
    cudaMemsetAsync(h->status.p, 0, 4, h->stream);
    return fail(h, OS1_E_CAPACITY, m);
}

// the four launches of one extraction over `nframes` frames whose level 0 is already on the device
// (frames [f0, f0+nframes) of the handle's per-frame buffers; kp/desc/counts are already offset by the caller)
int run_pipeline(os1_orb* h, const uint8_t* lvl0, long long lvl0Stride, int lvl0Pitch, int nframes, os1_kp_dev* kp, uint8_t* desc,
                 int cap, int* counts, int f0 = 0, cudaStream_t st = nullptr)
{
    if (!st) st = h->stream;
    const Geometry& g = h->geom;
    DevPtrs d;
    const long long F = f0;
    d.lvl0 = lvl0 + F * lvl0Stride; d.lvl0FrameStride = lvl0Stride; d.lvl0Pitch = lvl0Pitch;
    d.slab = (uint8_t*)h->slab.p + F * g.slabBytes;
    d.rtab = (const int16_t*)h->rtab.p; d.ptab = (const uint32_t*)h->ptab.p; d.gtab = (const int2*)h->gtab.p;
    d.cellCnt = (uint32_t*)h->cellCnt.p + F * g.cellsPerFrame;
    d.stage = (uint32_t*)h->stage.p + F * g.candPerFrame; d.cand = (uint32_t*)h->cand.p + F * g.candPerFrame;
    d.code = (uint32_t*)h->code.p + F * g.candPerFrame; d.node = (uint16_t*)h->node.p + F * g.candPerFrame;
    d.candCount = (uint32_t*)h->candCount.p + F * g.nlevels;
    d.lvKp = (uint32_t*)h->lvKp.p + F * g.kpCapPerFrame; d.lvCount = (uint32_t*)h->lvCount.p + F * g.nlevels;
    d.status = (int*)h->status.p;
    if (f0 == 0) { h->lastLvl0 = lvl0; h->lastLvl0Stride = lvl0Stride; h->lastLvl0Pitch = lvl0Pitch; }
    h->lastFrames = f0 + nframes;
    PyramidPlan* const plan = (nframes >= 4 && h->planBig) ? h->planBig : h->plan;
    h->lastLaunches = (f0 == 0 ? 0 : h->lastLaunches) + os1_pyramid_launch_count(plan, g, nframes) + 3;
    const bool t = h->timing && f0 == 0;
    if (t) CK(cudaEventRecord(h->ev[0], st), "cudaEventRecord");
    CK(os1_launch_pyramid(plan, g, d, nframes, st), "pyramid kernel");
    if (t) CK(cudaEventRecord(h->ev[1], st), "cudaEventRecord");
    CK(os1_launch_fast(g, d, nframes, st), "FAST kernel");
    if (t) CK(cudaEventRecord(h->ev[2], st), "cudaEventRecord");
    CK(os1_launch_quadtree(g, d, nframes, h->maxQuota, h->maxCells, st), "quadtree kernel");
    if (t) CK(cudaEventRecord(h->ev[3], st), "cudaEventRecord");
    CK(os1_launch_describe(g, d, nframes, kp, desc, cap, counts, st), "describe kernel");
    if (t) CK(cudaEventRecord(h->ev[4], st), "cudaEventRecord");
    return OS1_OK;
}

bool tma_ok(const void* p, size_t pitch, size_t frameStride)
{
    return ((uintptr_t)p % 16 == 0) && (pitch % 16 == 0) && (frameStride % 16 == 0);
}

} // namespace

extern "C" {

int os1_orb_create(os1_orb** out, int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device)
{
    os1_orb* h = nullptr;
    if (!out) return fail(nullptr, OS1_E_INVALID, "os1_orb_create: out is NULL");
    *out = nullptr;
    if (nfeatures < 1 || nlevels < 1 || !(scaleFactor > 1.0f)) return fail(nullptr, OS1_E_INVALID, "os1_orb_create: need nfeatures >= 1, nlevels >= 1, scaleFactor > 1");
    if (nlevels > OS1_MAX_LEVELS) return fail(nullptr, OS1_E_UNSUPPORTED, "os1_orb_create: more than 12 pyramid levels");
    if (iniThFAST < 1 || minThFAST < 1 || iniThFAST > 254 || minThFAST > 254)
        return fail(nullptr, OS1_E_UNSUPPORTED, "os1_orb_create: FAST thresholds must lie in 1..254");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(nullptr, OS1_E_CUDA, std::string("os1_orb_create: no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    }
    if (device < 0 || device >= ndev) return fail(nullptr, OS1_E_INVALID, "os1_orb_create: bad device ordinal");
    h = new os1_orb();
    h->nfeatures = nfeatures; h->nlevels = nlevels; h->iniTh = iniThFAST; h->minTh = minThFAST;
    h->scaleFactor = scaleFactor;
    h->device = device;
    build_tables(h);
    if (!os1_quadtree_fits(h->maxQuota, 0, 1)) {      // (the image size adds its cell table to this: checked again per geometry)
        delete h;
        return fail(nullptr, OS1_E_UNSUPPORTED, "os1_orb_create: too many features in one pyramid level (the quadtree state of a level must fit one CTA's shared memory: about 2300)");
    }
    DeviceScope _ds(device);
    if ((e = _ds.err) != cudaSuccess || (e = cudaStreamCreateWithFlags(&h->ownStream, cudaStreamNonBlocking)) != cudaSuccess) {
        delete h;
        return cuda_fail(nullptr, e, "os1_orb_create");
    }
    h->stream = h->ownStream;
    *out = h;
    return OS1_OK;
}

int os1_orb_destroy(os1_orb* h)
{
    if (!h) return OS1_OK;
    DeviceScope _ds(h->device);
    cudaStreamSynchronize(h->stream);
    if (h->latExec) cudaGraphExecDestroy(h->latExec);
    if (h->plan) os1_pyramid_plan_destroy(h->plan);
    if (h->planBig) os1_pyramid_plan_destroy(h->planBig);
    DevBuf* bufs[] = { &h->rtab, &h->ptab, &h->gtab, &h->slab, &h->cellCnt, &h->stage, &h->cand, &h->code, &h->node, &h->candCount,
                       &h->lvKp, &h->lvCount, &h->status, &h->outKp, &h->outDesc, &h->outCount };
    for (DevBuf* b : bufs) b->release();
    if (h->hostOut) cudaFreeHost(h->hostOut);
    if (h->hostStatus) cudaFreeHost(h->hostStatus);
    for (int i = 0; i < 5; i++) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    for (cudaEvent_t e : h->chunkEv) cudaEventDestroy(e);
    for (int i = 0; i < 3; i++) {
        if (h->altStream[i]) cudaStreamDestroy(h->altStream[i]);
        if (h->altEv[i]) cudaEventDestroy(h->altEv[i]);
    }
    if (h->inStream) cudaStreamDestroy(h->inStream);
    if (h->outStream) cudaStreamDestroy(h->outStream);
    if (h->ownStream) cudaStreamDestroy(h->ownStream);
    delete h;
    return OS1_OK;
}

const char* os1_last_error(const os1_orb* h) { return h ? h->err.c_str() : t_err.c_str(); }

int os1_orb_get_levels(const os1_orb* h) { return h ? h->nlevels : 0; }
float os1_orb_get_scale_factor(const os1_orb* h) { return h ? (float)h->scaleFactor : 0.f; }

int os1_orb_get_tables(const os1_orb* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2)
{
    if (!h) return fail(nullptr, OS1_E_INVALID, "NULL handle");
    for (int i = 0; i < h->nlevels; i++) {
        if (scale) scale[i] = h->scale[i];
        if (inv_scale) inv_scale[i] = h->inv[i];
        if (sigma2) sigma2[i] = h->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = h->invsigma2[i];
    }
    return OS1_OK;
}

int os1_orb_get_features_per_level(const os1_orb* h, int* quota)
{
    if (!h || !quota) return fail(nullptr, OS1_E_INVALID, "NULL argument");
    for (int i = 0; i < h->nlevels; i++) quota[i] = h->quota[i];
    return OS1_OK;
}

int os1_orb_max_keypoints(const os1_orb* h)
{
    if (!h) return 0;
    // independent of the resolution: quota + 2 per level, or the 4*nIni nodes of the first pass (nIni <= 4 for
    // aspect ratios up to 4.5:1)
    int n = 0;
    for (int l = 0; l < h->nlevels; l++) n += std::max(h->quota[l] + 2, 16) + 2;
    return n;
}

int os1_orb_reserve(os1_orb* h, int rows, int cols, int max_frames)
{
    if (!h) return fail(nullptr, OS1_E_INVALID, "NULL handle");
    if (rows <= 0 || cols <= 0 || max_frames <= 0) return fail(h, OS1_E_INVALID, "os1_orb_reserve: bad size");
    DeviceScope _ds(h->device); CK(_ds.err, "cudaSetDevice");
    return ensure(h, rows, cols, max_frames);
}

int os1_orb_set_stream(os1_orb* h, void* cuda_stream)
{
    if (!h) return fail(nullptr, OS1_E_INVALID, "NULL handle");
    h->stream = cuda_stream ? (cudaStream_t)cuda_stream : h->ownStream;
    return OS1_OK;
}

int os1_orb_launches_per_call(const os1_orb* h) { return h ? h->lastLaunches : 4; }

int os1_orb_set_timing(os1_orb* h, int enable)
{
    if (!h) return fail(nullptr, OS1_E_INVALID, "NULL handle");
    DeviceScope _ds(h->device); CK(_ds.err, "cudaSetDevice");
    if (enable && !h->ev[0])
        for (int i = 0; i < 5; i++) CK(cudaEventCreate(&h->ev[i]), "cudaEventCreate");
    h->timing = enable != 0;
    return OS1_OK;
}

int os1_orb_get_stage_times(os1_orb* h, float* ms4)
{
    if (!h || !ms4) return fail(h, OS1_E_INVALID, "os1_orb_get_stage_times: NULL argument");
    if (!h->timing || !h->ev[0]) return fail(h, OS1_E_INVALID, "os1_orb_get_stage_times: timing is off");
    DeviceScope _ds(h->device); CK(_ds.err, "cudaSetDevice");
    CK(cudaEventSynchronize(h->ev[4]), "cudaEventSynchronize");
    for (int i = 0; i < 4; i++) CK(cudaEventElapsedTime(&ms4[i], h->ev[i], h->ev[i + 1]), "cudaEventElapsedTime");
    return OS1_OK;
}

int os1_orb_synchronize(os1_orb* h)
{
    if (!h) return fail(nullptr, OS1_E_INVALID, "NULL handle");
    DeviceScope _ds(h->device); CK(_ds.err, "cudaSetDevice");
    if (!h->status.p) { CK(cudaStreamSynchronize(h->stream), "cudaStreamSynchronize"); return OS1_OK; }
    CK(cudaMemcpyAsync(h->hostStatus, h->status.p, 4, cudaMemcpyDeviceToHost, h->stream), "read status");
    CK(cudaStreamSynchronize(h->stream), "cudaStreamSynchronize");
    return check_status(h, h->hostStatus[0]);
}

int os1_orb_extract_batch(os1_orb* h, const uint8_t* frames, int nframes, int rows, int cols, size_t step, size_t frame_stride,
                          int frames_on_device, os1_keypoint* kp, uint8_t* desc, int cap_per_frame, int* counts, int out_on_device)
{
    if (!h) return fail(nullptr, OS1_E_INVALID, "NULL handle");
    if (nframes <= 0 || !frames || rows <= 0 || cols <= 0) return OS1_OK;       // empty input: outputs untouched (:910-911)
    if (!kp || !desc || !counts || cap_per_frame <= 0) return fail(h, OS1_E_INVALID, "os1_orb_extract_batch: NULL output or zero capacity");
    if (step < (size_t)cols || (nframes > 1 && frame_stride < step * (size_t)(rows - 1) + cols))
        return fail(h, OS1_E_INVALID, "os1_orb_extract_batch: step/frame_stride smaller than the image");
    DeviceScope _ds(h->device); CK(_ds.err, "cudaSetDevice");
    int rc = ensure(h, rows, cols, nframes);
    if (rc) return rc;
    const Geometry& g = h->geom;
    const LevelGeom& L0 = g.lv[0];
    uint8_t* slab = (uint8_t*)h->slab.p;

    const int cap = g.kpCapPerFrame;
    const bool direct = frames_on_device && tma_ok(frames, step, frame_stride);      // read level 0 in place
    const cudaMemcpyKind inKind = frames_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    os1_kp_dev* dKp = out_on_device ? (os1_kp_dev*)kp : (os1_kp_dev*)h->outKp.p;
    uint8_t* dDesc = out_on_device ? desc : (uint8_t*)h->outDesc.p;
    int* dCnt = out_on_device ? counts : (int*)h->outCount.p;
    const int dCap = out_on_device ? cap_per_frame : cap;

    // stage frames [f0, f0+n) into the slab's level 0 (host frames: this IS the upload; odd device layouts: a D2D copy)
    auto stage_in = [&](int f0, int n, cudaStream_t st) -> cudaError_t {
        const uint8_t* src = frames + (size_t)f0 * frame_stride;
        uint8_t* dst = slab + (long long)f0 * g.slabBytes + L0.off;
        if ((size_t)L0.pitch == step) {
            // same row pitch on both sides: a frame is one contiguous run, the batch a 2-D copy of long rows
            // (a single frame needs no frame stride: whatever the caller passed there is not used as a pitch)
            const size_t run = (size_t)(rows - 1) * step + cols;
            if (n == 1) return cudaMemcpyAsync(dst, src, run, inKind, st);
            return cudaMemcpy2DAsync(dst, (size_t)g.slabBytes, src, frame_stride, run, n, inKind, st);
        }
        if (n > 1 && frame_stride % step == 0 && g.slabBytes % L0.pitch == 0) {
            cudaMemcpy3DParms p;
            memset(&p, 0, sizeof(p));
            p.srcPtr = make_cudaPitchedPtr((void*)src, step, cols, frame_stride / step);
            p.dstPtr = make_cudaPitchedPtr(dst, L0.pitch, cols, g.slabBytes / L0.pitch);
            p.extent = make_cudaExtent(cols, rows, n);
            p.kind = inKind;
            return cudaMemcpy3DAsync(&p, st);
        }
        for (int f = 0; f < n; f++) {
            cudaError_t e = cudaMemcpy2DAsync(dst + (long long)f * g.slabBytes, L0.pitch, src + (size_t)f * frame_stride, step, cols, rows, inKind, st);
            if (e != cudaSuccess) return e;
        }
        return cudaSuccess;
    };
    const uint8_t* lvl0 = direct ? frames : slab + L0.off;
    const long long lvl0Stride = direct ? (long long)frame_stride : g.slabBytes;
    const int lvl0Pitch = direct ? (int)step : L0.pitch;

    // ---- everything on the device: one launch sequence on the handle's stream, nothing to wait for ---------------------
    if (out_on_device && frames_on_device) {
        if (!direct) CK(stage_in(0, nframes, h->stream), "stage frames");
        return run_pipeline(h, lvl0, lvl0Stride, lvl0Pitch, nframes, dKp, dDesc, dCap, dCnt);
    }

    // ---- one frame from the host: the latency path --------------------------------------------------------------------------
    if (nframes == 1 && !out_on_device) {
        if (!direct) CK(stage_in(0, 1, h->stream), "upload frame");
        uint8_t* sb = (uint8_t*)h->status.p;                       // [status | count | pad to 64 | keypoints | descriptors]
        const size_t slabBytes = 64 + (size_t)cap * (sizeof(os1_kp_dev) + 32);
        uint8_t* hb = (uint8_t*)h->hostOut;
        os1_kp_dev* hk = (os1_kp_dev*)(hb + 64);
        uint8_t* hd = hb + 64 + (size_t)cap * sizeof(os1_kp_dev);
        int* hc = (int*)(hb + 4);
        static const bool useGraph = !(getenv("OS1_GRAPH") && atoi(getenv("OS1_GRAPH")) == 0);
        const bool graphOk = useGraph && !h->latDisabled && !frames_on_device && !h->timing;
        bool done = false;
        if (graphOk && h->latExec) {
            CK(cudaGraphLaunch(h->latExec, h->stream), "cudaGraphLaunch");
            h->lastFrames = 1; h->lastLvl0 = lvl0; h->lastLvl0Stride = lvl0Stride; h->lastLvl0Pitch = lvl0Pitch;    // (what run_pipeline records)
            done = true;
        } else if (graphOk && ++h->latCalls >= 3) {
            // (the first calls run plainly: lazily initialised tables and function attributes must exist before a capture)
            cudaGraph_t gph = nullptr;
            if (cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
                const int rcc = run_pipeline(h, lvl0, lvl0Stride, lvl0Pitch, 1, (os1_kp_dev*)(sb + 64), sb + 64 + (size_t)cap * sizeof(os1_kp_dev), cap, (int*)(sb + 4));
                cudaError_t ce = rcc ? cudaErrorUnknown : cudaMemcpyAsync(hb, sb, slabBytes, cudaMemcpyDeviceToHost, h->stream);
                cudaError_t ee = cudaStreamEndCapture(h->stream, &gph);
                if (ce == cudaSuccess && ee == cudaSuccess && gph && cudaGraphInstantiate(&h->latExec, gph, 0) == cudaSuccess) {
                    CK(cudaGraphLaunch(h->latExec, h->stream), "cudaGraphLaunch");
                    done = true;
                } else {
                    cudaGetLastError();
                    h->latExec = nullptr; h->latDisabled = true;          // this driver / stream cannot capture the sequence: plain launches from now on
                }
                if (gph) cudaGraphDestroy(gph);
            } else { cudaGetLastError(); h->latDisabled = true; }
        }
        if (!done) {
            rc = run_pipeline(h, lvl0, lvl0Stride, lvl0Pitch, 1, (os1_kp_dev*)(sb + 64), sb + 64 + (size_t)cap * sizeof(os1_kp_dev), cap, (int*)(sb + 4));
            if (rc) return rc;
            // one round trip, one copy: the status word, the count and the (small, fixed-size) output slabs travel together
            CK(cudaMemcpyAsync(hb, sb, slabBytes, cudaMemcpyDeviceToHost, h->stream), "download outputs");
        }
        CK(cudaStreamSynchronize(h->stream), "cudaStreamSynchronize");
        h->hostStatus[0] = *(int*)hb;
        rc = check_status(h, h->hostStatus[0]);
        if (rc) return rc;
        const int n = std::min(hc[0], cap);
        if (n > cap_per_frame) return fail(h, OS1_E_CAPACITY, "os1_orb_extract: caller capacity smaller than the keypoint count (see os1_orb_max_keypoints)");
        if (n > 0) {
            memcpy(kp, hk, (size_t)n * sizeof(os1_keypoint));
            memcpy(desc, hd, (size_t)n * 32);
        }
        counts[0] = n;
        return OS1_OK;
    }

    // ---- batches with host buffers: chunks flow through upload -> kernels -> download on three streams, so the PCIe
    // copies of neighbouring chunks hide behind the kernels of the current one ----------------------------------------------
    // (tuning knobs are read once per process)
    static const int envChunk = getenv("OS1_CHUNK") ? std::max(1, atoi(getenv("OS1_CHUNK"))) : 0;
    static const int envRamp0 = getenv("OS1_RAMP0") ? std::max(1, atoi(getenv("OS1_RAMP0"))) : 0;
    static const int envRampPct = getenv("OS1_RAMPPCT") ? std::max(100, atoi(getenv("OS1_RAMPPCT"))) : 0;
    // chunk = about 8.5 MB of frames (28 frames at 640x480, 9 at 1280x720, 4 at 1920x1080; measured optimum at all three: small
    // enough that little work is left when the last upload ends, large enough that the kernels of a chunk fill the GPU)
    const int chunk = envChunk ? envChunk : std::max(2, std::min(64, (int)lround(8.5e6 / ((double)rows * cols))));
    // chunk boundaries: a geometric ramp.  The upload of chunk c+1 hides behind the kernels of chunk c only while it is
    // not longer than them, so chunks grow by less than the kernel-time / upload-time ratio per frame (1.31 at 720p over
    // PCIe 5 x16); the kernels then start after the upload of a few frames and never starve again
    const int ramp0 = envRamp0 ? envRamp0 : std::max(1, chunk / 8), rampPct = envRampPct ? envRampPct : 125;
    std::vector<int> cuts(1, 0);
    for (int f = 0, sz = std::min(ramp0, chunk); f < nframes;) {
        f = std::min(nframes, f + sz);
        cuts.push_back(f);
        sz = std::min(chunk, std::max(sz + 1, sz * rampPct / 100));
    }
    const int nchunks = (int)cuts.size() - 1;
    if (!h->inStream) CK(cudaStreamCreateWithFlags(&h->inStream, cudaStreamNonBlocking), "cudaStreamCreate");
    if (!h->outStream) CK(cudaStreamCreateWithFlags(&h->outStream, cudaStreamNonBlocking), "cudaStreamCreate");
    // OS1_TRACE=1: the chunk events carry timestamps and the call prints, per chunk, when its upload, its kernels and its
    // download finished (developer aid for the pipeline's timeline; off by default)
    static const bool trace = getenv("OS1_TRACE") && atoi(getenv("OS1_TRACE")) != 0;
    while ((int)h->chunkEv.size() < 3 * nchunks + 2) {
        cudaEvent_t e;
        CK(cudaEventCreateWithFlags(&e, trace ? cudaEventDefault : cudaEventDisableTiming), "cudaEventCreate");
        h->chunkEv.push_back(e);
    }
    // the copy streams must not run ahead of work already queued on the compute stream (buffer reuse across calls)
    CK(cudaEventRecord(h->chunkEv[2 * nchunks], h->stream), "cudaEventRecord");
    CK(cudaStreamWaitEvent(h->inStream, h->chunkEv[2 * nchunks], 0), "cudaStreamWaitEvent");
    CK(cudaStreamWaitEvent(h->outStream, h->chunkEv[2 * nchunks], 0), "cudaStreamWaitEvent");
    // odd chunks run on a second compute stream (their scratch is disjoint): the partial last waves of one chunk's
    // kernels are filled by the other chunk's
    static const int nAlt = getenv("OS1_ALT") ? std::min(3, std::max(0, atoi(getenv("OS1_ALT")))) : 2;
    const bool alt = nAlt > 0 && nchunks > 1;
    if (alt) {
        for (int i = 0; i < nAlt; i++) {
            if (!h->altStream[i]) CK(cudaStreamCreateWithFlags(&h->altStream[i], cudaStreamNonBlocking), "cudaStreamCreate");
            if (!h->altEv[i]) CK(cudaEventCreateWithFlags(&h->altEv[i], cudaEventDisableTiming), "cudaEventCreate");
            CK(cudaStreamWaitEvent(h->altStream[i], h->chunkEv[2 * nchunks], 0), "cudaStreamWaitEvent");
        }
    }
    const size_t kpRow = (size_t)dCap * sizeof(os1_kp_dev), dRow = (size_t)dCap * 32;
    const int copyCap = std::min(dCap, cap_per_frame);
    // a failure half way through leaves copies queued that write into the caller's buffers: wait for them before returning
    auto drain = [&]() {
        cudaStreamSynchronize(h->inStream); cudaStreamSynchronize(h->outStream); cudaStreamSynchronize(h->stream);
        for (int i = 0; i < 3; i++) if (h->altStream[i]) cudaStreamSynchronize(h->altStream[i]);
        cudaGetLastError();
    };
    for (int c = 0; c < nchunks; c++) {
        const int f0 = cuts[c], n = cuts[c + 1] - cuts[c];
        cudaStream_t cs = (alt && c % (nAlt + 1)) ? h->altStream[c % (nAlt + 1) - 1] : h->stream;
        if (!direct) {
            CKD(stage_in(f0, n, frames_on_device ? cs : h->inStream), "upload frames");
            if (!frames_on_device) {
                CKD(cudaEventRecord(h->chunkEv[2 * c], h->inStream), "cudaEventRecord");
                CKD(cudaStreamWaitEvent(cs, h->chunkEv[2 * c], 0), "cudaStreamWaitEvent");
            }
        }
        rc = run_pipeline(h, lvl0, lvl0Stride, lvl0Pitch, n, dKp + (size_t)f0 * dCap, dDesc + (size_t)f0 * dRow, dCap, dCnt + f0, f0, cs);
        if (rc) { drain(); return rc; }
        if (!out_on_device) {
            CKD(cudaEventRecord(h->chunkEv[2 * c + 1], cs), "cudaEventRecord");
            CKD(cudaStreamWaitEvent(h->outStream, h->chunkEv[2 * c + 1], 0), "cudaStreamWaitEvent");
            // straight into the caller's slabs (row = frame; pitches differ when the caller's capacity does)
            CKD(cudaMemcpy2DAsync(kp + (size_t)f0 * cap_per_frame, (size_t)cap_per_frame * sizeof(os1_keypoint), dKp + (size_t)f0 * dCap, kpRow,
                                 (size_t)copyCap * sizeof(os1_keypoint), n, cudaMemcpyDeviceToHost, h->outStream), "download keypoints");
            CKD(cudaMemcpy2DAsync(desc + (size_t)f0 * cap_per_frame * 32, (size_t)cap_per_frame * 32, dDesc + (size_t)f0 * dRow, dRow,
                                 (size_t)copyCap * 32, n, cudaMemcpyDeviceToHost, h->outStream), "download descriptors");
            CKD(cudaMemcpyAsync(counts + f0, dCnt + f0, (size_t)n * 4, cudaMemcpyDeviceToHost, h->outStream), "download counts");
            if (trace) CKD(cudaEventRecord(h->chunkEv[2 * nchunks + 1 + c], h->outStream), "cudaEventRecord");
        }
    }
    if (alt) {       // join: everything queued on the second stream is ordered before what follows on the handle's stream
        for (int i = 0; i < nAlt; i++) {
            CKD(cudaEventRecord(h->altEv[i], h->altStream[i]), "cudaEventRecord");
            CKD(cudaStreamWaitEvent(h->stream, h->altEv[i], 0), "cudaStreamWaitEvent");
        }
    }
    CKD(cudaMemcpyAsync(h->hostStatus, h->status.p, 4, cudaMemcpyDeviceToHost, h->stream), "download status");
    CKD(cudaStreamSynchronize(h->stream), "cudaStreamSynchronize");
    if (!out_on_device) CKD(cudaStreamSynchronize(h->outStream), "cudaStreamSynchronize");
    if (trace && !out_on_device && !frames_on_device && !direct) {
        fprintf(stderr, "[os1 trace] %d frames in %d chunks; ms after the call's start: chunk(frames) upload / kernels / download\n", nframes, nchunks);
        for (int c = 0; c < nchunks; c++) {
            float a = 0, b = 0, d = 0;
            cudaEventElapsedTime(&a, h->chunkEv[2 * nchunks], h->chunkEv[2 * c]);
            cudaEventElapsedTime(&b, h->chunkEv[2 * nchunks], h->chunkEv[2 * c + 1]);
            cudaEventElapsedTime(&d, h->chunkEv[2 * nchunks], h->chunkEv[2 * nchunks + 1 + c]);
            fprintf(stderr, "[os1 trace]  %2d(%2d) %7.3f / %7.3f / %7.3f\n", c, cuts[c + 1] - cuts[c], a, b, d);
        }
        cudaGetLastError();
    }
    rc = check_status(h, h->hostStatus[0]);
    if (rc) return rc;
    if (!out_on_device)
        for (int f = 0; f < nframes; f++)
            if (counts[f] > cap_per_frame) return fail(h, OS1_E_CAPACITY, "os1_orb_extract_batch: caller capacity smaller than the keypoint count (see os1_orb_max_keypoints)");
    return OS1_OK;
}

int os1_orb_extract(os1_orb* h, const uint8_t* image, int rows, int cols, size_t step, os1_keypoint* kp, int kp_cap, uint8_t* desc, int* n)
{
    if (!h) return fail(nullptr, OS1_E_INVALID, "NULL handle");
    if (!image || rows <= 0 || cols <= 0) return OS1_OK;            // if(_image.empty()) return;  (:910-911)
    if (!n) return fail(h, OS1_E_INVALID, "os1_orb_extract: n is NULL");
    return os1_orb_extract_batch(h, image, 1, rows, cols, step, step * (size_t)rows, 0, kp, desc, kp_cap, n, 0);
}

// ---- stage accessors ------------------------------------------------------------------------------------------
int os1_orb_level_size(const os1_orb* h, int level, int* w, int* hh)
{
    if (!h || !h->haveGeom || level < 0 || level >= h->nlevels) return fail(nullptr, OS1_E_INVALID, "os1_orb_level_size: no geometry yet or bad level");
    if (w) *w = h->geom.lv[level].w;
    if (hh) *hh = h->geom.lv[level].h;
    return OS1_OK;
}

int os1_orb_get_pyramid_level(os1_orb* h, int frame, int level, uint8_t* out_host)
{
    if (!h || !h->haveGeom || level < 0 || level >= h->nlevels || frame < 0 || frame >= h->lastFrames || !out_host)
        return fail(h, OS1_E_INVALID, "os1_orb_get_pyramid_level: bad argument");
    DeviceScope _ds(h->device); CK(_ds.err, "cudaSetDevice");
    const LevelGeom& L = h->geom.lv[level];
    const uint8_t* src; size_t pitch;
    if (level == 0) { src = h->lastLvl0 + (long long)frame * h->lastLvl0Stride; pitch = h->lastLvl0Pitch; }
    else { src = (uint8_t*)h->slab.p + (long long)frame * h->geom.slabBytes + L.off; pitch = L.pitch; }
    CK(cudaStreamSynchronize(h->stream), "cudaStreamSynchronize");
    CK(cudaMemcpy2D(out_host, L.w, src, pitch, L.w, L.h, cudaMemcpyDeviceToHost), "download level");
    return OS1_OK;
}

static int fetch_packed(os1_orb* h, const uint32_t* dev, int n, int32_t* xys, int addBorder)
{
    std::vector<uint32_t> tmp(std::max(n, 1));
    if (n > 0) CK(cudaMemcpy(tmp.data(), dev, (size_t)n * 4, cudaMemcpyDeviceToHost), "download");
    for (int i = 0; i < n; i++) {
        xys[3 * i] = OS1_X(tmp[i]) + addBorder; xys[3 * i + 1] = OS1_Y(tmp[i]) + addBorder; xys[3 * i + 2] = OS1_S(tmp[i]);
    }
    return OS1_OK;
}

int os1_orb_get_candidates(os1_orb* h, int frame, int level, int32_t* xys, int cap, int* n)
{
    if (!h || !h->haveGeom || level < 0 || level >= h->nlevels || frame < 0 || frame >= h->lastFrames || !n)
        return fail(h, OS1_E_INVALID, "os1_orb_get_candidates: bad argument");
    DeviceScope _ds(h->device); CK(_ds.err, "cudaSetDevice");
    CK(cudaStreamSynchronize(h->stream), "cudaStreamSynchronize");
    uint32_t cnt = 0;
    CK(cudaMemcpy(&cnt, (uint32_t*)h->candCount.p + (size_t)frame * h->nlevels + level, 4, cudaMemcpyDeviceToHost), "download count");
    *n = (int)cnt;
    if (!xys) return OS1_OK;
    if ((int)cnt > cap) return fail(h, OS1_E_CAPACITY, "os1_orb_get_candidates: cap too small");
    return fetch_packed(h, (uint32_t*)h->cand.p + (size_t)frame * h->geom.candPerFrame + h->geom.lv[level].candOff, (int)cnt, xys, 0);
}

int os1_orb_get_level_keypoints(os1_orb* h, int frame, int level, int32_t* xys, int cap, int* n)
{
    if (!h || !h->haveGeom || level < 0 || level >= h->nlevels || frame < 0 || frame >= h->lastFrames || !n)
        return fail(h, OS1_E_INVALID, "os1_orb_get_level_keypoints: bad argument");
    DeviceScope _ds(h->device); CK(_ds.err, "cudaSetDevice");
    CK(cudaStreamSynchronize(h->stream), "cudaStreamSynchronize");
    uint32_t cnt = 0;
    CK(cudaMemcpy(&cnt, (uint32_t*)h->lvCount.p + (size_t)frame * h->nlevels + level, 4, cudaMemcpyDeviceToHost), "download count");
    *n = (int)cnt;
    if (!xys) return OS1_OK;
    if ((int)cnt > cap) return fail(h, OS1_E_CAPACITY, "os1_orb_get_level_keypoints: cap too small");
    return fetch_packed(h, (uint32_t*)h->lvKp.p + (size_t)frame * h->geom.kpCapPerFrame + h->geom.lv[level].kpOff, (int)cnt, xys, OS1_BORDER);
}

// ---- Hamming ----------------------------------------------------------------------------------------------------
int os1_descriptor_distance(const uint8_t* a, const uint8_t* b)
{
    uint32_t pa[8], pb[8];
    memcpy(pa, a, 32); memcpy(pb, b, 32);
    int d = 0;
    for (int i = 0; i < 8; i++) d += __builtin_popcount(pa[i] ^ pb[i]);
    return d;
}

static int sm_count(int device)
{
    static std::mutex mu;
    static int cache[64];
    std::lock_guard<std::mutex> lk(mu);
    if (device < 0 || device >= 64) return 148;
    if (!cache[device]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0) n = 148;
        cache[device] = n;
    }
    return cache[device];
}

// ---- scratch pool of os1_hamming_knn2 ----------------------------------------------------------------------------------
// A workspace is reused at once by a call on the stream that used it last (stream order keeps the two launches apart) and by
// any other stream once its last launch has finished; otherwise another workspace is made.  Tickets are zeroed when a
// workspace is created and are handed back at zero by every launch.
namespace {
struct HammingWs {
    int device = -1;
    uint2* partials = nullptr; size_t nPartials = 0;
    unsigned* tickets = nullptr; int nTickets = 0;
    cudaEvent_t done = nullptr;
    cudaStream_t lastStream = nullptr;
    bool used = false;
};
std::mutex g_hwsMu;
std::vector<HammingWs*> g_hws;

cudaError_t hws_acquire(int device, cudaStream_t s, size_t nPartials, int nTickets, HammingWs** out)
{
    HammingWs* ws = nullptr;
    for (HammingWs* w : g_hws) {
        if (w->device != device) continue;
        if (!w->used || w->lastStream == s || cudaEventQuery(w->done) == cudaSuccess) { ws = w; break; }
        cudaGetLastError();       // (cudaErrorNotReady is not an error)
    }
    if (!ws) {
        ws = new HammingWs();
        ws->device = device;
        cudaError_t e = cudaEventCreateWithFlags(&ws->done, cudaEventDisableTiming);
        if (e != cudaSuccess) { delete ws; return e; }
        g_hws.push_back(ws);
    }
    if (ws->nPartials < nPartials || ws->nTickets < nTickets) {
        if (ws->used) { cudaError_t e = cudaEventSynchronize(ws->done); if (e != cudaSuccess) return e; }
        const size_t np = std::max(nPartials, (size_t)64 * 1024);
        const int nt = std::max(nTickets, 256);
        if (ws->partials) cudaFree(ws->partials);
        if (ws->tickets) cudaFree(ws->tickets);
        ws->partials = nullptr; ws->tickets = nullptr; ws->nPartials = 0; ws->nTickets = 0;
        cudaError_t e = cudaMalloc((void**)&ws->partials, np * sizeof(uint2));
        if (e == cudaSuccess) e = cudaMalloc((void**)&ws->tickets, (size_t)nt * sizeof(unsigned));
        if (e == cudaSuccess) e = cudaMemset(ws->tickets, 0, (size_t)nt * sizeof(unsigned));
        if (e == cudaSuccess) e = cudaStreamSynchronize(0);
        if (e != cudaSuccess) return e;
        ws->nPartials = np; ws->nTickets = nt;
    }
    ws->used = true; ws->lastStream = s;
    *out = ws;
    return cudaSuccess;
}
} // namespace

int os1_hamming_knn2(const uint8_t* A, int nA, const uint8_t* B, int nB, int32_t* best_idx, int32_t* best_dist, int32_t* second_dist,
                     int on_device, int device, void* cuda_stream)
{
    os1_orb* h = nullptr;
    if (nA < 0 || nB < 0 || (nA > 0 && (!A || !best_idx || !best_dist || !second_dist)) || (nB > 0 && !B))
        return fail(nullptr, OS1_E_INVALID, "os1_hamming_knn2: bad argument");
    if (nA == 0) return OS1_OK;
    DeviceScope _ds(device); CK(_ds.err, "cudaSetDevice (no CPU fallback)");
    cudaStream_t s = (cudaStream_t)cuda_stream;
    const int nSplit = os1_hamming_splits(nA, nB, sm_count(device));
    // scratch (per-split partials + merge tickets) comes from a small per-process pool: a steady-state call allocates nothing
    std::lock_guard<std::mutex> lk(g_hwsMu);
    HammingWs* ws = nullptr;
    cudaError_t e = hws_acquire(device, s, (size_t)nSplit * nA, os1_hamming_row_blocks(nA), &ws);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "os1_hamming_knn2: workspace");
    if (on_device) {
        e = os1_launch_hamming_knn2_splits(A, nA, B, nB, best_idx, best_dist, second_dist, (uint64_t*)ws->partials, nSplit, s, ws->tickets);
        if (e == cudaSuccess) e = cudaEventRecord(ws->done, s);
        if (e != cudaSuccess) return cuda_fail(nullptr, e, "hamming kernel");
        return OS1_OK;
    }
    uint8_t *dA = nullptr, *dB = nullptr; int* dOut = nullptr;
    e = cudaMallocAsync((void**)&dA, (size_t)nA * 32, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dB, (size_t)std::max(nB, 1) * 32, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dOut, (size_t)nA * 12, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dA, A, (size_t)nA * 32, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess && nB > 0) e = cudaMemcpyAsync(dB, B, (size_t)nB * 32, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = os1_launch_hamming_knn2_splits(dA, nA, dB, nB, dOut, dOut + nA, dOut + 2 * nA, (uint64_t*)ws->partials, nSplit, s, ws->tickets);
    if (e == cudaSuccess) e = cudaEventRecord(ws->done, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(best_idx, dOut, (size_t)nA * 4, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(best_dist, dOut + nA, (size_t)nA * 4, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(second_dist, dOut + 2 * nA, (size_t)nA * 4, cudaMemcpyDeviceToHost, s);
    if (dA) cudaFreeAsync(dA, s);
    if (dB) cudaFreeAsync(dB, s);
    if (dOut) cudaFreeAsync(dOut, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "os1_hamming_knn2");
    return OS1_OK;
}

int os1_hamming_window(const uint8_t* A, int nA, const uint8_t* B, int nB, const int32_t* ptr, const int32_t* cand, int32_t* best_idx,
                       int32_t* best_dist, int32_t* second_dist, int on_device, int device, void* cuda_stream)
{
    os1_orb* h = nullptr;
    if (nA < 0 || nB < 0 || (nA > 0 && (!A || !ptr || !best_idx || !best_dist || !second_dist)))
        return fail(nullptr, OS1_E_INVALID, "os1_hamming_window: bad argument");
    if (nA == 0) return OS1_OK;
    DeviceScope _ds(device); CK(_ds.err, "cudaSetDevice (no CPU fallback)");
    cudaStream_t s = (cudaStream_t)cuda_stream;
    if (on_device) {
        cudaError_t e = os1_launch_hamming_window(A, nA, B, ptr, cand, best_idx, best_dist, second_dist, s);
        if (e != cudaSuccess) return cuda_fail(nullptr, e, "hamming window kernel");
        return OS1_OK;
    }
    const int total = ptr[nA];
    if (total < 0) return fail(nullptr, OS1_E_INVALID, "os1_hamming_window: ptr is not a prefix sum");
    for (int i = 0; i < nA; i++)
        if (ptr[i] > ptr[i + 1]) return fail(nullptr, OS1_E_INVALID, "os1_hamming_window: ptr is not monotone");
    for (int k = 0; k < total; k++)
        if (cand[k] < 0 || cand[k] >= nB) return fail(nullptr, OS1_E_INVALID, "os1_hamming_window: candidate index out of range");
    uint8_t *dA = nullptr, *dB = nullptr; int *dOut = nullptr, *dPtr = nullptr, *dCand = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&dA, (size_t)nA * 32, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dB, (size_t)std::max(nB, 1) * 32, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dOut, (size_t)nA * 12, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dPtr, (size_t)(nA + 1) * 4, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dCand, (size_t)std::max(total, 1) * 4, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dA, A, (size_t)nA * 32, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess && nB > 0) e = cudaMemcpyAsync(dB, B, (size_t)nB * 32, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dPtr, ptr, (size_t)(nA + 1) * 4, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess && total > 0) e = cudaMemcpyAsync(dCand, cand, (size_t)total * 4, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = os1_launch_hamming_window(dA, nA, dB, dPtr, dCand, dOut, dOut + nA, dOut + 2 * nA, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(best_idx, dOut, (size_t)nA * 4, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(best_dist, dOut + nA, (size_t)nA * 4, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(second_dist, dOut + 2 * nA, (size_t)nA * 4, cudaMemcpyDeviceToHost, s);
    void* frees[] = { dA, dB, dOut, dPtr, dCand };
    for (void* p : frees) if (p) cudaFreeAsync(p, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return cuda_fail(nullptr, e, "os1_hamming_window");
    return OS1_OK;
}

} // extern "C"
