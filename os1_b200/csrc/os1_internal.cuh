// os1_b200 — internal declarations shared by the CUDA translation units (sm_100a only).
//
// Geometry mirrors the reference extractor (paths relative to the os1 tree):
//   level sizes            src/ORBextractor.cc:976
//   FAST cell grid         src/ORBextractor.cc:807-844
//   quadtree roots/quotas  src/ORBextractor.cc:467-478, :575-595
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>

#define OS1_MAX_LEVELS 12
#define OS1_EDGE 19          // EDGE_THRESHOLD  (ORBextractor.cc:80)
#define OS1_BORDER 16        // minBorderX = EDGE_THRESHOLD-3 (:807)
#define OS1_PATH_BITS 12     // quadtree depth representable in a path code (level sides <= 4096 px)
#define FAST_G 8             // FAST cells per CTA (fast.cu)

// packed candidate / kept keypoint: x | y<<12 | score<<24  (x,y relative to (OS1_BORDER,OS1_BORDER))
#define OS1_PACK_XYS(x, y, s) ((uint32_t)(x) | ((uint32_t)(y) << 12) | ((uint32_t)(s) << 24))
#define OS1_X(p) ((int)((p) & 0xFFFu))
#define OS1_Y(p) ((int)(((p) >> 12) & 0xFFFu))
#define OS1_S(p) ((int)((p) >> 24))

struct LevelGeom {
    int w, h;               // level image size
    int pitch;              // bytes per row inside the pyramid slab (multiple of 16)
    long long off;          // byte offset of the level inside one frame's slab
    // FAST cell grid
    int nCols, nRows, wCell, hCell;
    int maxBX, maxBY;       // maxBorderX/Y = w-16, h-16
    int cellOff;            // first cell of this level in per-frame cell arrays
    int nCells;
    int cellCap;            // staging slots per cell
    long long candOff;      // first slot of this level in the per-frame staging / cand / code / node arrays
    int candCap;            // nCells * cellCap
    int groupOff;           // first FAST work group (CTA) of this level
    int groupsPerRow;       // CTAs per cell row
    // quadtree
    int W, H;               // maxBorder - minBorder
    int nIni, quota;
    int xsOff, ysOff;       // offsets (u32) of the spread path tables of this level
    int kpOff, kpCap;       // kept-keypoint slots of this level inside a frame
    // resize tables for this level as DESTINATION (level >= 1): entries {ofs, c0, c1}
    int rxOff, ryOff;
    float scale;            // mvScaleFactor[level]
    float kpSize;           // (float)(int)(31 * scale)
};

struct Geometry {
    int nlevels;
    int rows, cols;
    int iniTh, minTh;
    int cellsPerFrame;
    int groupsPerFrame;
    int kpCapPerFrame;          // sum of kpCap
    long long candPerFrame;     // sum of candCap
    long long slabBytes;        // pyramid slab bytes per frame
    LevelGeom lv[OS1_MAX_LEVELS];
};

// per-call device pointers
struct DevPtrs {
    const uint8_t* lvl0;        // level-0 pixels (user buffer or slab)
    long long lvl0FrameStride;  // bytes between frames of level 0
    int lvl0Pitch;
    uint8_t* slab;              // pyramid slabs (levels >= 1; level 0 too when staged)
    const int16_t* rtab;        // resize tables, 4 x int16 per entry {ofs, c0, c1, pad}
    const uint32_t* ptab;       // quadtree path tables
    const int2* gtab;           // FAST work groups of one frame: {level | cell row << 8, first cell column}
    uint32_t* cellCnt;          // [frame][cell]
    uint32_t* stage;            // [frame][candPerFrame]
    uint32_t* cand;             // [frame][candPerFrame]  ordered candidates
    uint32_t* code;             // [frame][candPerFrame]
    uint16_t* node;             // [frame][candPerFrame]
    uint32_t* candCount;        // [frame][level]
    uint32_t* lvKp;             // [frame][kpCapPerFrame] packed kept keypoints, level-major
    uint32_t* lvCount;          // [frame][level]
    int* status;                // device-side error flags (bit field), one int
};

#define OS1_ERR_NODE_OVERFLOW 1
#define OS1_ERR_DEPTH 2
#define OS1_ERR_KP_OVERFLOW 4
#define OS1_ERR_CELL_OVERFLOW 8

// cv::KeyPoint layout (28 bytes): pt.x pt.y size angle response octave class_id
struct os1_kp_dev { float x, y, size, angle, response; int octave, class_id; };

// Makes `device` current for the lifetime of the object and restores the caller's current device afterwards: the library
// never leaves a host thread pointed at a different GPU than it came in with.
struct DeviceScope {
    int prev = -1;
    cudaError_t err;
    explicit DeviceScope(int device)
    {
        err = cudaGetDevice(&prev);
        if (err != cudaSuccess) { prev = -1; cudaGetLastError(); }
        if (prev == device) { prev = -1; err = cudaSuccess; }
        else err = cudaSetDevice(device);
    }
    ~DeviceScope() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceScope(const DeviceScope&) = delete;
    DeviceScope& operator=(const DeviceScope&) = delete;
};

// ---- launchers (one per kernel family) -----------------------------------------------------------
struct PyramidPlan;   // host-side tile tables (pyramid.cu)
PyramidPlan* os1_pyramid_plan_create(const Geometry& g, int tileW, int tileH);
void os1_pyramid_plan_destroy(PyramidPlan*);
bool os1_pyramid_plan_ok(const PyramidPlan*);     // level-0 footprint fits a TMA box and the tile chain fits shared memory
cudaError_t os1_launch_pyramid(PyramidPlan*, const Geometry& g, const DevPtrs& d, int nframes, cudaStream_t s);
int os1_pyramid_launch_count(const PyramidPlan*, const Geometry& g, int nframes);      // kernels the call above launches
cudaError_t os1_launch_fast(const Geometry& g, const DevPtrs& d, int nframes, cudaStream_t s);
bool os1_quadtree_fits(int maxQuota, int maxCells, int maxIni);       // the per-level state fits one CTA's shared memory
cudaError_t os1_launch_quadtree(const Geometry& g, const DevPtrs& d, int nframes, int maxQuota, int maxCells, cudaStream_t s);
cudaError_t os1_describe_init();      // per-device constant table; call with the device current, outside any capture
cudaError_t os1_launch_describe(const Geometry& g, const DevPtrs& d, int nframes, os1_kp_dev* kp, uint8_t* desc, int capPerFrame,
                                int* counts, cudaStream_t s);
// Hamming: nSplit partitions of B; scratch holds nSplit * nA uint2 partials
int os1_hamming_splits(int nA, int nB, int sms);
int os1_hamming_row_blocks(int nA);      // merge tickets one launch needs (zero-initialised unsigned each)
cudaError_t os1_launch_hamming_knn2_splits(const uint8_t* A, int nA, const uint8_t* B, int nB, int* bestIdx, int* bestDist,
                                           int* secondDist, uint64_t* scratch, int nSplit, cudaStream_t s, unsigned* counters = nullptr);
cudaError_t os1_launch_hamming_window(const uint8_t* A, int nA, const uint8_t* B, const int* ptr, const int* cand, int* bestIdx,
                                      int* bestDist, int* secondDist, cudaStream_t s);
cudaError_t os1_launch_hamming_csr(const uint8_t* A, int nA, const uint8_t* B, const int* ptr, const int* cand, uint16_t* dist, cudaStream_t s,
                                   int cap = 0x7fffffff);
void os1_internal_set_error(const char* msg);
// Frame grid (framegrid.cu): Frame's 64 x 48 grid over the undistorted image bounds and GetFeaturesInArea on it
#define OS1_GRID_COLS 64
#define OS1_GRID_ROWS 48
#define OS1_GRID_CELLS (OS1_GRID_COLS * OS1_GRID_ROWS)
#define OS1_GRID_MAX_KP 16384
struct GridGeom { float minX, minY, invW, invH; };
cudaError_t os1_launch_grid_build(const os1_kp_dev* kp, int n, GridGeom G, int* cellStart, int* gridIdx, cudaStream_t s);
cudaError_t os1_launch_area_csr(const os1_kp_dev* kp, const int* cellStart, const int* gridIdx, GridGeom G, const float* qx, const float* qy,
                                const float* qr, const int* qmin, const int* qmax, const uint8_t* qvalid, int nq, int* cnt, int* ptr, int* cand,
                                int candCap, cudaStream_t s);
// Raises a kernel's dynamic shared-memory limit to the device's opt-in maximum, once per (kernel, device) and under a
// lock, so that handles with different tile sizes — and host threads launching at the same time — never lower or race
// on a limit another launch relies on (the attribute belongs to the function, not to a handle).
cudaError_t os1_optin_smem(const void* kernel);
cudaError_t os1_launch_hamming_pairs(const uint8_t* desc, const int* counts, int cap, const int* pairs, int npairs, int* bestIdx, int* bestDist,
                                     int* secondDist, int* accepted, float nnratio, int thLow, cudaStream_t s);
