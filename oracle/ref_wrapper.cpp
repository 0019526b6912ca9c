// TEST INFRASTRUCTURE — NOT PRODUCT CODE (see oracle/cv_exact.h header).
//
// C wrapper around the UNMODIFIED reference extractor: /root/reference/src/ORBextractor.cc is compiled
// by path (oracle/Makefile) against oracle/cvshim; nothing from the reference is copied into this repo.
// Built twice into oracle/_ref/:
//   libos1_ref.so        glibc allocator  — the reference as it really behaves; the timed CPU baseline
//   libos1_ref_canon.so  -DREF_BUMP_ALLOC — operator new is a never-reusing bump arena, so std::list node
//                        addresses grow with creation order and the pointer tie-break of
//                        std::sort(pair<int,ExtractorNode*>) (ORBextractor.cc:716) becomes the declared
//                        canonical rule "later-created node sorts higher" (SURVEY.md Appendix C).
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <new>
#include <vector>
#include "ORBextractor.h"   // the reference's own header, found through -I/root/reference/include

#ifdef REF_BUMP_ALLOC
#include <sys/mman.h>
namespace {
const size_t ARENA_BYTES = (size_t)24 << 30;    // virtual reservation; pages are touched on demand
char* g_arena = nullptr;
size_t g_top = 0;
bool g_active = false;
inline bool in_arena(void* p) { return g_arena && (char*)p >= g_arena && (char*)p < g_arena + ARENA_BYTES; }
void arena_begin()
{
    if (!g_arena) {
        void* p = mmap(nullptr, ARENA_BYTES, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (p == MAP_FAILED) { perror("oracle bump arena"); abort(); }
        g_arena = (char*)p;
    }
    g_top = 0;
    g_active = true;
}
void arena_end() { g_active = false; }
}
void* operator new(size_t n)
{
    if (g_active) {
        size_t a = (g_top + 15) & ~(size_t)15;
        if (a + n > ARENA_BYTES) { fprintf(stderr, "oracle bump arena exhausted\n"); abort(); }
        g_top = a + n;
        return g_arena + a;
    }
    void* p = malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
void* operator new[](size_t n) { return operator new(n); }
void operator delete(void* p) noexcept { if (p && !in_arena(p)) free(p); }
void operator delete[](void* p) noexcept { operator delete(p); }
void operator delete(void* p, size_t) noexcept { operator delete(p); }
void operator delete[](void* p, size_t) noexcept { operator delete(p); }
#else
namespace { void arena_begin() {} void arena_end() {} }
#endif

namespace {

struct Probe : public ORB_SLAM2::ORBextractor {
    Probe(int n, float s, int l, int a, int b) : ORB_SLAM2::ORBextractor(n, s, l, a, b) {}
    using ORB_SLAM2::ORBextractor::DistributeOctTree;
    using ORB_SLAM2::ORBextractor::mnFeaturesPerLevel;
    using ORB_SLAM2::ORBextractor::umax;
    using ORB_SLAM2::ORBextractor::nlevels;
};

// FAST-call recorder: candidates per level in the order the reference accumulates vToDistributeKeys
// (malloc/realloc storage on purpose: it must not live in the bump arena, which is reset between calls)
struct Cand { int x, y, score; };
struct CandBuf {
    Cand* p; size_t n, cap;
    void push(const Cand& c)
    {
        if (n == cap) { cap = cap ? cap * 2 : 4096; p = (Cand*)realloc(p, cap * sizeof(Cand)); }
        p[n++] = c;
    }
};
const int MAX_LEVELS = 64;
thread_local CandBuf g_cands[MAX_LEVELS];
thread_local long long g_fast_px[2] = { 0, 0 };      // pixels handed to cv::FAST at the first / the retry threshold of the last recorded call
thread_local int g_fast_th0 = -1;
thread_local int g_cand_levels = 0;
thread_local Probe* t_probe = nullptr;

void on_fast(const cv::Mat& img, int threshold, const std::vector<cv::KeyPoint>& kps)
{
    if (!t_probe) return;
    if (g_fast_th0 < 0) g_fast_th0 = threshold;        // the first call of an extraction uses iniThFAST (ORBextractor.cc:848)
    g_fast_px[threshold == g_fast_th0 ? 0 : 1] += (long long)img.rows * img.cols;
    if (kps.empty()) return;
    for (int l = 0; l < t_probe->nlevels && l < MAX_LEVELS; l++) {
        const cv::Mat& L = t_probe->mvImagePyramid[l];
        const uint8_t* lo = L.data;
        const uint8_t* hi = L.data + (size_t)(L.rows - 1) * L.step + L.cols;
        if (img.data >= lo && img.data < hi) {
            size_t off = img.data - lo;
            int oy = (int)(off / L.step), ox = (int)(off % L.step);
            for (size_t i = 0; i < kps.size(); i++)   // coordinates relative to (minBorderX, minBorderY) = (16,16)
                g_cands[l].push(Cand{ (int)kps[i].pt.x + ox - 16, (int)kps[i].pt.y + oy - 16, (int)kps[i].response });
            return;
        }
    }
}

} // namespace

extern "C" {

void* ref_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh)
{
    return new Probe(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
void ref_destroy(void* h) { delete (Probe*)h; }

int ref_levels(void* h) { return ((Probe*)h)->GetLevels(); }
void ref_tables(void* h, float* scale, float* inv, float* sigma2, float* invsigma2, int* quota, int* umax16)
{
    Probe* p = (Probe*)h;
    std::vector<float> a = p->GetScaleFactors(), b = p->GetInverseScaleFactors(), c = p->GetScaleSigmaSquares(),
                       d = p->GetInverseScaleSigmaSquares();
    for (int i = 0; i < p->nlevels; i++) {
        scale[i] = a[i]; inv[i] = b[i]; sigma2[i] = c[i]; invsigma2[i] = d[i];
        quota[i] = p->mnFeaturesPerLevel[i];
    }
    for (int i = 0; i < 16; i++) umax16[i] = p->umax[i];
}

// Full operator().  kp rows: x, y, size, angle, response, octave (6 floats).  Returns the keypoint count
// (which may exceed `cap`; only `cap` rows are written).  If record != 0 the FAST candidates are kept for
// ref_candidates().

int ref_extract(void* h, const uint8_t* img, int rows, int cols, int step, float* kp, int cap, uint8_t* desc, int record)
{
    Probe* p = (Probe*)h;
    int n = 0;
    if (record) {
        g_fast_px[0] = g_fast_px[1] = 0; g_fast_th0 = -1;
        for (int l = 0; l < MAX_LEVELS; l++) g_cands[l].n = 0;
        g_cand_levels = p->nlevels;
        t_probe = p;
        cv::fast_hook() = on_fast;
    } else {
        cv::fast_hook() = nullptr;
    }
    arena_begin();
    {
        cv::Mat image(rows, cols, CV_8UC1, (void*)img, (size_t)step);
        std::vector<cv::KeyPoint> kps;
        cv::Mat d;
        (*p)(image, cv::Mat(), kps, d);
        n = (int)kps.size();
        for (int i = 0; i < n && i < cap; i++) {
            kp[i * 6 + 0] = kps[i].pt.x; kp[i * 6 + 1] = kps[i].pt.y; kp[i * 6 + 2] = kps[i].size;
            kp[i * 6 + 3] = kps[i].angle; kp[i * 6 + 4] = kps[i].response; kp[i * 6 + 5] = (float)kps[i].octave;
            if (desc) memcpy(desc + (size_t)i * 32, d.ptr(i), 32);
        }
    }
    arena_end();
    t_probe = nullptr;
    cv::fast_hook() = nullptr;
    return n;
}

int ref_pyramid_size(void* h, int level, int* w, int* hh)
{
    Probe* p = (Probe*)h;
    const cv::Mat& L = p->mvImagePyramid[level];
    *w = L.cols; *hh = L.rows;
    return 0;
}
// copy level (without the 19-px border if border==0, else the whole bordered buffer)
int ref_pyramid_level(void* h, int level, uint8_t* out, int border)
{
    Probe* p = (Probe*)h;
    const cv::Mat& L = p->mvImagePyramid[level];
    int b = border ? 19 : 0;
    int W = L.cols + 2 * b;
    for (int y = -b; y < L.rows + b; y++)
        memcpy(out + (size_t)(y + b) * W, L.data + (ptrdiff_t)y * (ptrdiff_t)L.step - b, W);
    return 0;
}

// pixels the last recorded extraction handed to cv::FAST: [0] at iniThFAST (:848), [1] at the minThFAST retry (:854)
void ref_fast_area(long long* px2) { px2[0] = g_fast_px[0]; px2[1] = g_fast_px[1]; }

int ref_candidate_count(int level) { return level < g_cand_levels ? (int)g_cands[level].n : 0; }
int ref_candidates(int level, int* xys /* n x 3 */)
{
    if (level >= g_cand_levels) return 0;
    const CandBuf& v = g_cands[level];
    for (size_t i = 0; i < v.n; i++) { xys[i * 3] = v.p[i].x; xys[i * 3 + 1] = v.p[i].y; xys[i * 3 + 2] = v.p[i].score; }
    return (int)v.n;
}

// DistributeOctTree alone (ORBextractor.cc:571-795) on a caller-supplied candidate list.
int ref_distribute(void* h, const int* xys, int n, int minX, int maxX, int minY, int maxY, int N, int level, int* out_xys, int cap)
{
    Probe* p = (Probe*)h;
    int m = 0;
    arena_begin();
    {
        std::vector<cv::KeyPoint> in(n);
        for (int i = 0; i < n; i++)
            in[i] = cv::KeyPoint((float)xys[i * 3], (float)xys[i * 3 + 1], 7.f, -1, (float)xys[i * 3 + 2]);
        std::vector<cv::KeyPoint> out = p->DistributeOctTree(in, minX, maxX, minY, maxY, N, level);
        m = (int)out.size();
        for (int i = 0; i < m && i < cap; i++) {
            out_xys[i * 3] = (int)out[i].pt.x; out_xys[i * 3 + 1] = (int)out[i].pt.y; out_xys[i * 3 + 2] = (int)out[i].response;
        }
    }
    arena_end();
    return m;
}

} // extern "C"
