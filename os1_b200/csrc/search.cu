// Matcher selection loops on top of the batched distance kernel (include/os1_b200.h, "matcher" section).
//
// The three reference loops named by the north star walk their queries IN ORDER and carry state from one query to the
// next (claimed partners in SearchForInitialization, src/ORBmatcher.cc:439,:458-466; map points already attached to
// the frame in the two SearchByProjection variants, :1363-1365,:1381 and :65-67,:118).  That bookkeeping is inherently
// sequential and tiny; the arithmetic — one 256-bit Hamming distance per (query, candidate) — is not.  So every distance
// of the candidate lists is computed in one launch of hamming_csr_kernel, and the ordered pass below runs on the host
// over those distances, reproducing the reference's comparisons one for one.
#include "../../include/os1_b200.h"
#include "os1_internal.cuh"
#include <algorithm>
#include <climits>
#include <cmath>
#include <string>
#include <vector>

namespace {

const int TH_LOW = 50, TH_HIGH = 100, HISTO_LENGTH = 30;      // src/ORBmatcher.cc:35-37

// ComputeThreeMaxima : src/ORBmatcher.cc:1554-1595
void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int i = 0; i < L; i++) {
        const int s = (int)histo[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

int rot_bin(float a1, float a2)
{
    const float factor = 1.0f / HISTO_LENGTH;
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)roundf(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

int validate(int nA, int nB, const int32_t* ptr, const int32_t* cand, const char* who)
{
    if (nA < 0 || nB < 0 || (nA > 0 && !ptr)) { os1_internal_set_error((std::string(who) + ": bad argument").c_str()); return OS1_E_INVALID; }
    for (int i = 0; i < nA; i++)
        if (ptr[i] > ptr[i + 1] || ptr[i] < 0) { os1_internal_set_error((std::string(who) + ": ptr is not a prefix sum").c_str()); return OS1_E_INVALID; }
    const int total = nA > 0 ? ptr[nA] : 0;
    if (total > 0 && !cand) { os1_internal_set_error((std::string(who) + ": cand is NULL").c_str()); return OS1_E_INVALID; }
    for (int k = 0; k < total; k++)
        if (cand[k] < 0 || cand[k] >= nB) { os1_internal_set_error((std::string(who) + ": candidate index out of range").c_str()); return OS1_E_INVALID; }
    return OS1_OK;
}

} // namespace

extern "C" {

int os1_hamming_csr(const uint8_t* A, int nA, const uint8_t* B, int nB, const int32_t* ptr, const int32_t* cand, uint16_t* dist, int on_device,
                    int device, void* cuda_stream)
{
    if (nA <= 0) return OS1_OK;
    cudaStream_t s = (cudaStream_t)cuda_stream;
    DeviceScope _ds(device); cudaError_t e = _ds.err;
    if (e != cudaSuccess) { cudaGetLastError(); os1_internal_set_error("os1_hamming_csr: cudaSetDevice failed (no CPU fallback)"); return OS1_E_CUDA; }
    if (on_device) {
        e = os1_launch_hamming_csr(A, nA, B, ptr, cand, dist, s);
        if (e != cudaSuccess) { cudaGetLastError(); os1_internal_set_error("os1_hamming_csr: launch failed"); return OS1_E_CUDA; }
        return OS1_OK;
    }
    int rc = validate(nA, nB, ptr, cand, "os1_hamming_csr");
    if (rc) return rc;
    const int total = ptr[nA];
    if (total == 0) return OS1_OK;
    if (!A || !B || !dist) { os1_internal_set_error("os1_hamming_csr: NULL buffer"); return OS1_E_INVALID; }
    uint8_t *dA = nullptr, *dB = nullptr; int *dPtr = nullptr, *dCand = nullptr; uint16_t* dDist = nullptr;
    e = cudaMallocAsync((void**)&dA, (size_t)nA * 32, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dB, (size_t)nB * 32, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dPtr, (size_t)(nA + 1) * 4, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dCand, (size_t)total * 4, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dDist, (size_t)total * 2, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dA, A, (size_t)nA * 32, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dB, B, (size_t)nB * 32, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dPtr, ptr, (size_t)(nA + 1) * 4, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dCand, cand, (size_t)total * 4, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = os1_launch_hamming_csr(dA, nA, dB, dPtr, dCand, dDist, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dist, dDist, (size_t)total * 2, cudaMemcpyDeviceToHost, s);
    void* frees[] = { dA, dB, dPtr, dCand, dDist };
    for (void* p : frees) if (p) cudaFreeAsync(p, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) { cudaGetLastError(); os1_internal_set_error("os1_hamming_csr: CUDA failure"); return OS1_E_CUDA; }
    return OS1_OK;
}

int os1_hamming_knn2_pairs(const uint8_t* desc, const int32_t* counts, int cap, const int32_t* pairs, int npairs, int32_t* best_idx,
                           int32_t* best_dist, int32_t* second_dist, int32_t* accepted, float nnratio, int th_low, int device, void* cuda_stream)
{
    if (npairs < 0 || cap <= 0 || (npairs > 0 && (!desc || !counts || !pairs))) { os1_internal_set_error("os1_hamming_knn2_pairs: bad argument"); return OS1_E_INVALID; }
    if (npairs == 0) return OS1_OK;
    DeviceScope _ds(device); cudaError_t e = _ds.err;
    if (e == cudaSuccess) e = os1_launch_hamming_pairs(desc, counts, cap, pairs, npairs, best_idx, best_dist, second_dist, accepted, nnratio, th_low,
                                                      (cudaStream_t)cuda_stream);
    if (e != cudaSuccess) { cudaGetLastError(); os1_internal_set_error("os1_hamming_knn2_pairs: CUDA failure (no CPU fallback)"); return OS1_E_CUDA; }
    return OS1_OK;
}

int os1_search_for_initialization(const uint8_t* D1, int n1, const uint8_t* D2, int n2, const int32_t* ptr, const int32_t* cand, float nnratio,
                                  const float* angles1, const float* angles2, int32_t* matches12, int* nmatches, int device)
{
    if (!matches12 || !nmatches) { os1_internal_set_error("os1_search_for_initialization: NULL output"); return OS1_E_INVALID; }
    int rc = validate(n1, n2, ptr, cand, "os1_search_for_initialization");
    if (rc) return rc;
    const int total = n1 > 0 ? ptr[n1] : 0;
    std::vector<uint16_t> dist(std::max(total, 1));
    rc = os1_hamming_csr(D1, n1, D2, n2, ptr, cand, dist.data(), 0, device, nullptr);
    if (rc) return rc;
    // src/ORBmatcher.cc:400-515
    int nm = 0;
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    std::vector<int> matchedDist(std::max(n2, 1), INT_MAX), matches21(std::max(n2, 1), -1);
    std::vector<int> rotHist[HISTO_LENGTH];
    const bool checkOri = angles1 && angles2;
    for (int i1 = 0; i1 < n1; i1++) {
        if (ptr[i1] == ptr[i1 + 1]) continue;                              // level > 0 or empty window (:417,:422)
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int k = ptr[i1]; k < ptr[i1 + 1]; k++) {
            const int i2 = cand[k], d = dist[k];
            if (matchedDist[i2] <= d) continue;                             // :439
            if (d < bestDist) { bestDist2 = bestDist; bestDist = d; bestIdx2 = i2; }
            else if (d < bestDist2) bestDist2 = d;
        }
        if (bestDist <= TH_LOW && bestDist < (float)bestDist2 * nnratio) {   // :454-456
            if (matches21[bestIdx2] >= 0) { matches12[matches21[bestIdx2]] = -1; nm--; }
            matches12[i1] = bestIdx2; matches21[bestIdx2] = i1; matchedDist[bestIdx2] = bestDist;
            nm++;
            if (checkOri) rotHist[rot_bin(angles1[i1], angles2[bestIdx2])].push_back(i1);
        }
    }
    if (checkOri) {
        int ind1, ind2, ind3;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; nm--; }
        }
    }
    *nmatches = nm;
    return OS1_OK;
}

int os1_search_by_projection_frame(const uint8_t* Dq, int nq, const uint8_t* Dc, int nc, const int32_t* ptr, const int32_t* cand,
                                   uint8_t* taken, const float* angles_q, const float* angles_c, int32_t* match_of_cand, int* nmatches, int device)
{
    if (!match_of_cand || !nmatches || (nc > 0 && !taken)) { os1_internal_set_error("os1_search_by_projection_frame: NULL argument"); return OS1_E_INVALID; }
    int rc = validate(nq, nc, ptr, cand, "os1_search_by_projection_frame");
    if (rc) return rc;
    const int total = nq > 0 ? ptr[nq] : 0;
    std::vector<uint16_t> dist(std::max(total, 1));
    rc = os1_hamming_csr(Dq, nq, Dc, nc, ptr, cand, dist.data(), 0, device, nullptr);
    if (rc) return rc;
    // src/ORBmatcher.cc:1313-1419
    int nm = 0;
    for (int i = 0; i < nc; i++) match_of_cand[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const bool checkOri = angles_q && angles_c;
    for (int i = 0; i < nq; i++) {
        if (ptr[i] == ptr[i + 1]) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (int k = ptr[i]; k < ptr[i + 1]; k++) {
            const int i2 = cand[k];
            if (taken[i2]) continue;                                         // :1363-1365
            const int d = dist[k];
            if (d < bestDist) { bestDist = d; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {                                           // :1379
            match_of_cand[bestIdx2] = i; taken[bestIdx2] = 1;                // :1381 (the attached point has observations)
            nm++;
            if (checkOri) rotHist[rot_bin(angles_q[i], angles_c[bestIdx2])].push_back(bestIdx2);
        }
    }
    if (checkOri) {
        int ind1, ind2, ind3;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { match_of_cand[idx] = -1; taken[idx] = 0; nm--; }     // :1410-1416
        }
    }
    *nmatches = nm;
    return OS1_OK;
}

int os1_search_by_projection_points(const uint8_t* Dq, int nq, const uint8_t* Dc, int nc, const int32_t* ptr, const int32_t* cand,
                                    uint8_t* taken, const int32_t* octave_c, float nnratio, int32_t* match_of_cand, int* nmatches, int device)
{
    if (!match_of_cand || !nmatches || (nc > 0 && (!taken || !octave_c))) { os1_internal_set_error("os1_search_by_projection_points: NULL argument"); return OS1_E_INVALID; }
    int rc = validate(nq, nc, ptr, cand, "os1_search_by_projection_points");
    if (rc) return rc;
    const int total = nq > 0 ? ptr[nq] : 0;
    std::vector<uint16_t> dist(std::max(total, 1));
    rc = os1_hamming_csr(Dq, nq, Dc, nc, ptr, cand, dist.data(), 0, device, nullptr);
    if (rc) return rc;
    // src/ORBmatcher.cc:45-124
    int nm = 0;
    for (int i = 0; i < nc; i++) match_of_cand[i] = -1;
    for (int i = 0; i < nq; i++) {
        if (ptr[i] == ptr[i + 1]) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int k = ptr[i]; k < ptr[i + 1]; k++) {
            const int idx = cand[k];
            if (taken[idx]) continue;                                        // :83-85
            const int d = dist[k];
            if (d < bestDist) { bestDist2 = bestDist; bestDist = d; bestLevel2 = bestLevel; bestLevel = octave_c[idx]; bestIdx = idx; }
            else if (d < bestDist2) { bestLevel2 = octave_c[idx]; bestDist2 = d; }
        }
        if (bestDist <= TH_HIGH) {                                           // :111
            if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;     // :113-114
            match_of_cand[bestIdx] = i; taken[bestIdx] = 1;                  // :116
            nm++;
        }
    }
    *nmatches = nm;
    return OS1_OK;
}

} // extern "C"
