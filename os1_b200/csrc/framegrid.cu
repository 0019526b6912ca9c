// Frame-side candidate generation: Frame::AssignFeaturesToGrid / PosInGrid
// (src/Frame.cc:114-129, :264-274) and Frame::GetFeaturesInArea (src/Frame.cc:209-262) for a whole batch of queries,
// producing the CSR candidate lists the matcher kernels consume — in the reference's visiting order: grid column (ix)
// outer, grid row (iy) inner, and inside a cell ascending keypoint index (AssignFeaturesToGrid pushes in index order).
//
// Three small kernels: (1) one CTA sorts the keys (cell << 16 | index) in shared memory (bitonic) and writes the
// per-cell start offsets; (2) a thread per query walks its cell window and counts the keypoints passing the level and
// |dx| < r, |dy| < r tests; a block scan turns counts into ptr; (3) the same walk writes the indices.  All float
// expressions are spelled exactly as the reference evaluates them (sub, sub, mul; roundf / floorf / ceilf).
#include "../../include/os1_b200.h"
#include "os1_internal.cuh"
#include <algorithm>
#include <cstring>
#include <vector>

#define GRID_COLS 64        // FRAME_GRID_COLS  include/Frame.h:37
#define GRID_ROWS 48        // FRAME_GRID_ROWS  include/Frame.h:36
#define GRID_CELLS (GRID_COLS * GRID_ROWS)
#define FG_THREADS 1024
#define FG_MAX_KP 16384

namespace {

__global__ void __launch_bounds__(FG_THREADS) grid_build_kernel(const os1_kp_dev* kp, int n, GridGeom G, int n2, int* cellStart, int* gridIdx)
{
    extern __shared__ unsigned keys[];       // n2 (power of two >= n)
    const int tid = threadIdx.x;
    for (int i = tid; i < n2; i += FG_THREADS) {
        unsigned k = 0xFFFFFFFFu;
        if (i < n) {
            // PosInGrid: round((kp.pt.x-mnMinX)*mfGridElementWidthInv)   Frame.cc:266-267
            const int px = (int)roundf(__fmul_rn(__fsub_rn(kp[i].x, G.minX), G.invW));
            const int py = (int)roundf(__fmul_rn(__fsub_rn(kp[i].y, G.minY), G.invH));
            if (px >= 0 && px < GRID_COLS && py >= 0 && py < GRID_ROWS) k = ((unsigned)(px * GRID_ROWS + py) << 16) | (unsigned)i;
        }
        keys[i] = k;
    }
    __syncthreads();
    for (int k = 2; k <= n2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < n2; i += FG_THREADS) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const unsigned a = keys[i], b = keys[ixj];
                    if ((a > b) == ((i & k) == 0)) { keys[i] = b; keys[ixj] = a; }
                }
            }
            __syncthreads();
        }
    // cellStart[c] = first position whose cell >= c
    for (int i = tid; i < n2; i += FG_THREADS) {
        const unsigned k = keys[i];
        const int c = k == 0xFFFFFFFFu ? GRID_CELLS : (int)(k >> 16);
        const int cp = i == 0 ? -1 : (keys[i - 1] == 0xFFFFFFFFu ? GRID_CELLS : (int)(keys[i - 1] >> 16));
        for (int cc = cp + 1; cc <= c; cc++) cellStart[cc] = i;
        if (i == n2 - 1) for (int cc = c + 1; cc <= GRID_CELLS; cc++) cellStart[cc] = n2;
        if (i < n) gridIdx[i] = k == 0xFFFFFFFFu ? -1 : (int)(k & 0xFFFFu);
    }
}

// the body of GetFeaturesInArea (Frame.cc:209-262) for one query; WRITE = false counts, true stores
template <bool WRITE>
__device__ __forceinline__ int area_walk(const os1_kp_dev* kp, const int* cellStart, const int* gridIdx, const GridGeom& G, float x, float y, float r,
                                         int minLevel, int maxLevel, int* out)
{
    const int nMinCellX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, G.minX), r), G.invW)));
    if (nMinCellX >= GRID_COLS) return 0;
    const int nMaxCellX = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, G.minX), r), G.invW)));
    if (nMaxCellX < 0) return 0;
    const int nMinCellY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, G.minY), r), G.invH)));
    if (nMinCellY >= GRID_ROWS) return 0;
    const int nMaxCellY = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, G.minY), r), G.invH)));
    if (nMaxCellY < 0) return 0;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    int cnt = 0;
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
        // cells (ix, nMinCellY..nMaxCellY) are contiguous in the sorted order (cell id = ix * ROWS + iy)
        const int b = cellStart[ix * GRID_ROWS + nMinCellY], e = cellStart[ix * GRID_ROWS + nMaxCellY + 1];
        for (int p = b; p < e; p++) {
            const int i = gridIdx[p];
            const os1_kp_dev k = kp[i];
            if (bCheckLevels) {
                if (k.octave < minLevel) continue;
                if (maxLevel >= 0 && k.octave > maxLevel) continue;
            }
            const float distx = __fsub_rn(k.x, x), disty = __fsub_rn(k.y, y);
            if (fabsf(distx) < r && fabsf(disty) < r) {
                if (WRITE) out[cnt] = i;
                cnt++;
            }
        }
    }
    return cnt;
}

// qvalid (may be NULL): queries the caller's loop skips (`continue`) have qvalid[q] == 0 and get an empty list
__global__ void area_count_kernel(const os1_kp_dev* kp, const int* cellStart, const int* gridIdx, GridGeom G, const float* qx, const float* qy,
                                  const float* qr, const int* qmin, const int* qmax, const uint8_t* qvalid, int nq, int* cnt)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    if (qvalid && !qvalid[q]) { cnt[q] = 0; return; }
    cnt[q] = area_walk<false>(kp, cellStart, gridIdx, G, qx[q], qy[q], qr[q], qmin ? qmin[q] : -1, qmax ? qmax[q] : -1, nullptr);
}

__global__ void __launch_bounds__(FG_THREADS) area_scan_kernel(const int* cnt, int nq, int* ptr)
{
    __shared__ int warpSums[33];
    __shared__ int carry;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nq; base += FG_THREADS) {
        const int i = base + tid;
        const int v = i < nq ? cnt[i] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        if (lane == 31) warpSums[w] = inc;
        __syncthreads();
        if (w == 0) {
            int s = warpSums[lane], si = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, si, o); if (lane >= o) si += t; }
            warpSums[lane] = si - s;
            if (lane == 31) warpSums[32] = si;
        }
        __syncthreads();
        if (i < nq) ptr[i] = carry + warpSums[w] + inc - v;
        __syncthreads();
        if (tid == 0) carry += warpSums[32];
        __syncthreads();
    }
    if (tid == 0) ptr[nq] = carry;
}

__global__ void area_fill_kernel(const os1_kp_dev* kp, const int* cellStart, const int* gridIdx, GridGeom G, const float* qx, const float* qy,
                                 const float* qr, const int* qmin, const int* qmax, const uint8_t* qvalid, int nq, const int* ptr, int* cand, int candCap)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    if (qvalid && !qvalid[q]) return;
    if (ptr[q + 1] > candCap) return;        // the caller's list is too short: reported through ptr[nq] > cand_cap
    area_walk<true>(kp, cellStart, gridIdx, G, qx[q], qy[q], qr[q], qmin ? qmin[q] : -1, qmax ? qmax[q] : -1, cand + ptr[q]);
}

} // namespace

// grid of a frame's keypoints: cellStart[GRID_CELLS + 1], gridIdx[n]
cudaError_t os1_launch_grid_build(const os1_kp_dev* kp, int n, GridGeom G, int* cellStart, int* gridIdx, cudaStream_t s)
{
    int n2 = 1; while (n2 < n) n2 <<= 1;
    const size_t smem = (size_t)n2 * 4;
    if (smem > 48 * 1024) { cudaError_t e = os1_optin_smem((const void*)grid_build_kernel); if (e != cudaSuccess) return e; }
    grid_build_kernel<<<1, FG_THREADS, smem, s>>>(kp, n, G, n2, cellStart, gridIdx);
    return cudaGetLastError();
}

// candidate lists of nq queries against a built grid: cnt[nq] scratch, ptr[nq + 1], cand[candCap]
cudaError_t os1_launch_area_csr(const os1_kp_dev* kp, const int* cellStart, const int* gridIdx, GridGeom G, const float* qx, const float* qy,
                                const float* qr, const int* qmin, const int* qmax, const uint8_t* qvalid, int nq, int* cnt, int* ptr, int* cand,
                                int candCap, cudaStream_t s)
{
    if (nq <= 0) return cudaMemsetAsync(ptr, 0, 4, s);
    area_count_kernel<<<(nq + 127) / 128, 128, 0, s>>>(kp, cellStart, gridIdx, G, qx, qy, qr, qmin, qmax, qvalid, nq, cnt);
    area_scan_kernel<<<1, FG_THREADS, 0, s>>>(cnt, nq, ptr);
    area_fill_kernel<<<(nq + 127) / 128, 128, 0, s>>>(kp, cellStart, gridIdx, G, qx, qy, qr, qmin, qmax, qvalid, nq, ptr, cand, candCap);
    return cudaGetLastError();
}

namespace {

cudaError_t run_device(const os1_kp_dev* kp, int n, GridGeom G, const float* qx, const float* qy, const float* qr, const int* qmin, const int* qmax,
                       int nq, int* ptr, int* cand, int candCap, cudaStream_t s)
{
    int *cellStart = nullptr, *gridIdx = nullptr, *cnt = nullptr;
    cudaError_t e = cudaMallocAsync((void**)&cellStart, (GRID_CELLS + 1) * 4, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&gridIdx, (size_t)std::max(n, 1) * 4, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&cnt, (size_t)std::max(nq, 1) * 4, s);
    if (e == cudaSuccess) e = os1_launch_grid_build(kp, n, G, cellStart, gridIdx, s);
    if (e == cudaSuccess) e = os1_launch_area_csr(kp, cellStart, gridIdx, G, qx, qy, qr, qmin, qmax, nullptr, nq, cnt, ptr, cand, candCap, s);
    if (cellStart) cudaFreeAsync(cellStart, s);
    if (gridIdx) cudaFreeAsync(gridIdx, s);
    if (cnt) cudaFreeAsync(cnt, s);
    return e;
}

} // namespace

extern "C" int os1_features_in_area(const os1_keypoint* kp, int n, float min_x, float max_x, float min_y, float max_y, const float* qx,
                                    const float* qy, const float* qr, const int32_t* q_min_level, const int32_t* q_max_level, int nq,
                                    int32_t* ptr, int32_t* cand, int cand_cap, int on_device, int device, void* cuda_stream)
{
    if (n < 0 || nq < 0 || !ptr || (nq > 0 && (!qx || !qy || !qr)) || (n > 0 && !kp) || cand_cap < 0 || (cand_cap > 0 && !cand)) {
        os1_internal_set_error("os1_features_in_area: bad argument");
        return OS1_E_INVALID;
    }
    if (n > FG_MAX_KP) { os1_internal_set_error("os1_features_in_area: more than 16384 keypoints in a frame"); return OS1_E_UNSUPPORTED; }
    if (!(max_x > min_x) || !(max_y > min_y)) { os1_internal_set_error("os1_features_in_area: empty image bounds"); return OS1_E_INVALID; }
    GridGeom G;
    G.minX = min_x; G.minY = min_y;
    G.invW = (float)GRID_COLS / (max_x - min_x);          // mfGridElementWidthInv   Frame.cc:99
    G.invH = (float)GRID_ROWS / (max_y - min_y);
    cudaStream_t s = (cudaStream_t)cuda_stream;
    DeviceScope _ds(device); cudaError_t e = _ds.err;
    if (e != cudaSuccess) { cudaGetLastError(); os1_internal_set_error("os1_features_in_area: cudaSetDevice failed (no CPU fallback)"); return OS1_E_CUDA; }
    if (on_device) {
        e = run_device((const os1_kp_dev*)kp, n, G, qx, qy, qr, q_min_level, q_max_level, nq, ptr, cand, cand_cap, s);
        if (e != cudaSuccess) { cudaGetLastError(); os1_internal_set_error("os1_features_in_area: CUDA failure"); return OS1_E_CUDA; }
        return OS1_OK;
    }
    os1_kp_dev* dKp = nullptr; float* dQ = nullptr; int *dLv = nullptr, *dPtr = nullptr, *dCand = nullptr;
    e = cudaMallocAsync((void**)&dKp, (size_t)std::max(n, 1) * sizeof(os1_kp_dev), s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dQ, (size_t)std::max(nq, 1) * 12, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dLv, (size_t)std::max(nq, 1) * 8, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dPtr, (size_t)(nq + 1) * 4, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dCand, (size_t)std::max(cand_cap, 1) * 4, s);
    if (e == cudaSuccess && n > 0) e = cudaMemcpyAsync(dKp, kp, (size_t)n * sizeof(os1_kp_dev), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess && nq > 0) {
        e = cudaMemcpyAsync(dQ, qx, (size_t)nq * 4, cudaMemcpyHostToDevice, s);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dQ + nq, qy, (size_t)nq * 4, cudaMemcpyHostToDevice, s);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dQ + 2 * nq, qr, (size_t)nq * 4, cudaMemcpyHostToDevice, s);
        if (e == cudaSuccess && q_min_level) e = cudaMemcpyAsync(dLv, q_min_level, (size_t)nq * 4, cudaMemcpyHostToDevice, s);
        if (e == cudaSuccess && q_max_level) e = cudaMemcpyAsync(dLv + nq, q_max_level, (size_t)nq * 4, cudaMemcpyHostToDevice, s);
    }
    if (e == cudaSuccess)
        e = run_device(dKp, n, G, dQ, dQ + nq, dQ + 2 * nq, q_min_level ? dLv : nullptr, q_max_level ? dLv + nq : nullptr, nq, dPtr, dCand, cand_cap, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(ptr, dPtr, (size_t)(nq + 1) * 4, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    int rc = OS1_OK;
    if (e == cudaSuccess) {
        if (ptr[nq] > cand_cap) { os1_internal_set_error("os1_features_in_area: cand_cap smaller than the number of candidates (ptr[nq] holds the need)"); rc = OS1_E_CAPACITY; }
        else if (ptr[nq] > 0) e = cudaMemcpyAsync(cand, dCand, (size_t)ptr[nq] * 4, cudaMemcpyDeviceToHost, s);
    }
    void* frees[] = { dKp, dQ, dLv, dPtr, dCand };
    for (void* p : frees) if (p) cudaFreeAsync(p, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) { cudaGetLastError(); os1_internal_set_error("os1_features_in_area: CUDA failure"); return OS1_E_CUDA; }
    return rc;
}

// ---- Frame::UndistortKeyPoints (src/Frame.cc:286-320) ------------------------------------------------------------------------
// mode 0: cv::undistortPoints(pts, pts, K, dist, Mat(), K) — five fixed-point iterations of the inverse Brown model in double
// (the library is compiled with -fmad=false, so every product and sum rounds as the x86 build of OpenCV rounds it);
// mode 1: os1's equidistant fisheye, antidistorsionarProyeccionEquidistante (:355-384).
namespace {

struct CamParams { double fx, fy, cx, cy; double k[14]; float Kf[9]; int mode; };

__global__ void undistort_kernel(const os1_kp_dev* in, int n, CamParams C, os1_kp_dev* out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    os1_kp_dev kp = in[i];
    if (C.mode == 0) {
        const double ifx = 1. / C.fx, ify = 1. / C.fy;
        double x = kp.x, y = kp.y;
        x = (x - C.cx) * ifx; y = (y - C.cy) * ify;
        const double x0 = x, y0 = y;
        const double* k = C.k;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = x0; y = y0; break; }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        const double xx = C.fx * x + 0.0 * y + C.cx, yy = 0.0 * x + C.fy * y + C.cy, ww = 1. / (0.0 * x + 0.0 * y + 1.0);
        kp.x = (float)(xx * ww); kp.y = (float)(yy * ww);
    } else {
        const double pwx = ((double)kp.x - C.cx) / C.fx, pwy = ((double)kp.y - C.cy) / C.fy;
        const double theta_d = sqrt(pwx * pwx + pwy * pwy);
        const double scale = theta_d > 1e-8 ? tan(theta_d) / theta_d : 1.0;
        const double pux = pwx * scale, puy = pwy * scale;
        const double pr0 = C.Kf[0] * pux + C.Kf[1] * puy + C.Kf[2] * 1.0, pr1 = C.Kf[3] * pux + C.Kf[4] * puy + C.Kf[5] * 1.0;
        const double pr2 = C.Kf[6] * pux + C.Kf[7] * puy + C.Kf[8] * 1.0;
        kp.x = (float)(pr0 / pr2); kp.y = (float)(pr1 / pr2);
    }
    out[i] = kp;
}

} // namespace

extern "C" int os1_undistort_keypoints(const os1_keypoint* kp, int n, const float* K4, const float* dist, int ndist, int mode, os1_keypoint* kp_un,
                                       int on_device, int device, void* cuda_stream)
{
    if (n < 0 || !K4 || (n > 0 && (!kp || !kp_un)) || ndist < 0 || ndist > 14 || (ndist > 0 && !dist) || (mode != 0 && mode != 1)) {
        os1_internal_set_error("os1_undistort_keypoints: bad argument");
        return OS1_E_INVALID;
    }
    if (n == 0) return OS1_OK;
    cudaStream_t s = (cudaStream_t)cuda_stream;
    DeviceScope _ds(device); cudaError_t e = _ds.err;
    if (e != cudaSuccess) { cudaGetLastError(); os1_internal_set_error("os1_undistort_keypoints: cudaSetDevice failed (no CPU fallback)"); return OS1_E_CUDA; }
    const bool identity = mode == 0 && (ndist == 0 || dist[0] == 0.0f);        // Frame.cc:288-292
    CamParams C;
    C.fx = K4[0]; C.fy = K4[1]; C.cx = K4[2]; C.cy = K4[3]; C.mode = mode;
    for (int i = 0; i < 14; i++) C.k[i] = i < ndist ? (double)dist[i] : 0.0;
    const float Kf[9] = { K4[0], 0.f, K4[2], 0.f, K4[1], K4[3], 0.f, 0.f, 1.f };
    for (int i = 0; i < 9; i++) C.Kf[i] = Kf[i];
    const size_t bytes = (size_t)n * sizeof(os1_kp_dev);
    if (on_device) {
        if (identity) e = kp == kp_un ? cudaSuccess : cudaMemcpyAsync(kp_un, kp, bytes, cudaMemcpyDeviceToDevice, s);
        else { undistort_kernel<<<(n + 127) / 128, 128, 0, s>>>((const os1_kp_dev*)kp, n, C, (os1_kp_dev*)kp_un); e = cudaGetLastError(); }
        if (e != cudaSuccess) { cudaGetLastError(); os1_internal_set_error("os1_undistort_keypoints: CUDA failure"); return OS1_E_CUDA; }
        return OS1_OK;
    }
    if (identity) { if (kp != kp_un) memcpy(kp_un, kp, bytes); return OS1_OK; }
    os1_kp_dev* d = nullptr;
    e = cudaMallocAsync((void**)&d, bytes, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d, kp, bytes, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) { undistort_kernel<<<(n + 127) / 128, 128, 0, s>>>(d, n, C, d); e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaMemcpyAsync(kp_un, d, bytes, cudaMemcpyDeviceToHost, s);
    if (d) cudaFreeAsync(d, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) { cudaGetLastError(); os1_internal_set_error("os1_undistort_keypoints: CUDA failure"); return OS1_E_CUDA; }
    return OS1_OK;
}

// Frame::ComputeImageBounds (src/Frame.cc:322-353): the four image corners through the same undistortion
extern "C" int os1_image_bounds(int cols, int rows, const float* K4, const float* dist, int ndist, int mode, float* bounds4, int device)
{
    if (!bounds4 || !K4 || cols <= 0 || rows <= 0 || (mode != 0 && mode != 1) || ndist < 0 || ndist > 14 || (mode == 0 && ndist > 0 && !dist)) {
        os1_internal_set_error("os1_image_bounds: bad argument (NULL pointer, mode outside {0,1}, or ndist outside 0..14)");
        return OS1_E_INVALID;
    }
    if (mode == 0 && (ndist == 0 || dist[0] == 0.0f)) {
        bounds4[0] = 0.0f; bounds4[1] = (float)cols; bounds4[2] = 0.0f; bounds4[3] = (float)rows;
        return OS1_OK;
    }
    os1_keypoint c[4];
    memset(c, 0, sizeof(c));
    c[0].x = 0.0f; c[0].y = 0.0f; c[1].x = (float)cols; c[1].y = 0.0f; c[2].x = 0.0f; c[2].y = (float)rows; c[3].x = (float)cols; c[3].y = (float)rows;
    const int rc = os1_undistort_keypoints(c, 4, K4, dist, ndist, mode, c, 0, device, nullptr);
    if (rc) return rc;
    bounds4[0] = std::min(c[0].x, c[2].x); bounds4[1] = std::max(c[1].x, c[3].x);       // mnMinX, mnMaxX
    bounds4[2] = std::min(c[0].y, c[1].y); bounds4[3] = std::max(c[2].y, c[3].y);       // mnMinY, mnMaxY
    return OS1_OK;
}
