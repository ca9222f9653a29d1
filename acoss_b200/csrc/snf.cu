// snf.cu -- host side of the similarity-network-fusion front end: acoss_upload_aux_songs and
// acoss_snf_pairs.  Kernels in kernels_snf.cuh; the binarise + Qmax / Dmax tail is the existing EarlySNF tail
// (acoss_align_dense_resident, api.cu).
#include <algorithm>
#include <cmath>

#include "ctx.h"
#include "kernels_snf.cuh"

using namespace acoss;

namespace {

// get_oti (CRPUtils.py:109-136) on the host copy of the global chroma: numpy's pairwise summation order for 12 values
// (eight running sums combined pairwise, then the tail), separately rounded products, first maximum wins -- the same
// arithmetic as k_oti (kernels_prep.cuh).
int host_oti(const double* gi, const double* gj, int bins) {
    int best = 0;
    double best_v = 0.0;
    for (int s = 0; s < bins; ++s) {
        volatile double prod[kMaxBins];
        for (int k = 0; k < bins; ++k) {
            int src = k - s;
            if (src < 0) src += bins;
            prod[k] = gi[src] * gj[k];
        }
        volatile double res;
        if (bins < 8) {
            res = 0.0;
            for (int i = 0; i < bins; ++i) res = res + prod[i];
        } else {
            volatile double r[8];
            for (int k = 0; k < 8; ++k) r[k] = prod[k];
            int i = 8;
            for (; i < bins - (bins % 8); i += 8)
                for (int k = 0; k < 8; ++k) r[k] = r[k] + prod[i + k];
            volatile double a = r[0] + r[1], b = r[2] + r[3], c2 = r[4] + r[5], d2 = r[6] + r[7];
            volatile double ab = a + b, cd = c2 + d2;
            res = ab + cd;
            for (; i < bins; ++i) res = res + prod[i];
        }
        if (s == 0 || res > best_v) {
            best_v = res;
            best = s;
        }
    }
    return best;
}

template <class K, class... A>
int launch2d(acoss_ctx* c, K kfn, int nx, int ny, A... args) {
    if (nx <= 0 || ny <= 0) return ACOSS_OK;
#ifdef ACOSS_EMU
    emu::launch(dim3((unsigned)nx, (unsigned)ny), dim3((unsigned)kSnfThreads), 0, [&]() { kfn(args...); });
#else
    kfn<<<dim3((unsigned)nx, (unsigned)ny), kSnfThreads, 0, c->stream>>>(args...);
#endif
    c->launches++;
    CU(cudaGetLastError());
    return ACOSS_OK;
}

}  // namespace

extern "C" {

int acoss_upload_aux_songs(acoss_ctx* c, const void* frames, int dtype, int32_t bins) {
    if (!c) return fail(ACOSS_ERR_ARG, "null context");
    CU(cudaSetDevice(c->device));
    if (c->n_songs <= 0) return fail(ACOSS_ERR_STATE, "acoss_upload_songs has not been called");
    if (!frames) return fail(ACOSS_ERR_ARG, "null input");
    if (bins < 1 || bins > kMaxBins) return fail(ACOSS_ERR_ARG, "bins must be in 1..%d", kMaxBins);
    if (dtype != ACOSS_F32 && dtype != ACOSS_F64) return fail(ACOSS_ERR_ARG, "unknown dtype %d", dtype);
    const size_t n_el = (size_t)c->total_frames * bins;
    CU(c->frames2.reserve(std::max<size_t>(n_el, 1) * sizeof(double)));
    if (dtype == ACOSS_F64) {
        CU(cudaMemcpyAsync(c->frames2.p, frames, n_el * sizeof(double), cudaMemcpyDefault, c->stream));
        CU(cudaStreamSynchronize(c->stream));
    } else {
        // widen on the host side of the copy: float32 -> float64 through a temporary (this arena is uploaded once)
        std::vector<float> tmp(n_el);
        CU(cudaMemcpy(tmp.data(), frames, n_el * sizeof(float), cudaMemcpyDefault));
        std::vector<double> wide(tmp.begin(), tmp.end());
        CU(cudaMemcpy(c->frames2.p, wide.data(), n_el * sizeof(double), cudaMemcpyHostToDevice));
    }
    c->bins2 = bins;
    return ACOSS_OK;
}

int acoss_snf_pairs(acoss_ctx* c, const int32_t* pairs, int64_t P, int32_t m, double kappa, int32_t niters, uint32_t flags,
                    float* qmax_raw, float* dmax_raw, int32_t* oti_out, int32_t* shape_out, double* fused_out,
                    const int64_t* fused_offsets, uint32_t* crp_out, const int64_t* crp_offsets) {
    if (!c) return fail(ACOSS_ERR_ARG, "null context");
    CU(cudaSetDevice(c->device));
    if (c->n_songs <= 0) return fail(ACOSS_ERR_STATE, "acoss_upload_songs has not been called");
    if (c->bins2 <= 0) return fail(ACOSS_ERR_STATE, "acoss_upload_aux_songs has not been called (second feature kind)");
    if (P < 0 || (P > 0 && !pairs)) return fail(ACOSS_ERR_ARG, "bad pair list");
    if (m < 1 || m > 32) return fail(ACOSS_ERR_ARG, "m must be in 1..32");
    if (!(kappa > 0.0 && kappa < 1.0)) return fail(ACOSS_ERR_ARG, "kappa must be in (0, 1)");
    if (niters < 1) return fail(ACOSS_ERR_ARG, "niters must be >= 1");
    if ((fused_out && !fused_offsets) || (crp_out && !crp_offsets)) return fail(ACOSS_ERR_ARG, "an output needs its offsets");
    const bool no_oti = (flags & kFlagNoOti) != 0;
    if (!no_oti && c->h_gchroma.empty()) return fail(ACOSS_ERR_STATE, "OTI requested but no gchroma uploaded");
    const int64_t budget = (int64_t)6 << 30;  // bytes of n x n matrices per batch
    int64_t p0 = 0;
    while (p0 < P) {
        std::vector<SnfDesc> D;
        std::vector<int32_t> rows, cols;
        std::vector<int64_t> which;
        int64_t mats = 0, md = 0, sp = 0, dense = 0;
        int nmax = 0, Mmax = 0;
        int64_t p = p0;
        for (; p < P && (int)D.size() < 4096; ++p) {
            const int32_t i = pairs[2 * p], j = pairs[2 * p + 1];
            if (i < 0 || i >= c->n_songs || j < 0 || j >= c->n_songs)
                return fail(ACOSS_ERR_ARG, "pair %lld = (%d, %d) out of range (n_songs %d)", (long long)p, i, j, c->n_songs);
            const int M = c->h_len[i] - m + 1, N = c->h_len[j] - m + 1;
            if (shape_out) {
                shape_out[2 * p] = M;
                shape_out[2 * p + 1] = N;
            }
            const int oti = no_oti ? 0 : host_oti(&c->h_gchroma[(size_t)i * c->bins], &c->h_gchroma[(size_t)j * c->bins], c->bins);
            if (oti_out) oti_out[p] = oti;
            if (qmax_raw) qmax_raw[p] = 0.0f;
            if (dmax_raw) dmax_raw[p] = 0.0f;
            if (M <= 0 || N <= 0) continue;   // a song shorter than the window: no plot, scores 0 (as the path does)
            SnfDesc d;
            memset(&d, 0, sizeof(d));
            d.M = M;
            d.N = N;
            d.n = M + N;
            d.K = (int)(kappa * (double)(M + N));                       // EarlySNF.py:52
            d.k1 = (int)((double)d.K * (double)M / (double)(M + N));    // SimilarityFusion.py:126
            d.k2 = d.K - d.k1;
            if (d.n > kSnfMaxN) return fail(ACOSS_ERR_ARG, "pair %lld: M + N = %d exceeds %d", (long long)p, d.n, kSnfMaxN);
            if (d.K > kSnfMaxK || d.k1 < 1 || d.k2 < 1 || d.k1 + 1 >= M || d.k2 + 1 >= N || d.K >= d.n)
                return fail(ACOSS_ERR_ARG, "pair %lld (%d x %d): K = %d, k1 = %d, k2 = %d is outside what the fusion supports", (long long)p, M, N,
                            d.K, d.k1, d.k2);
            const int64_t nn = (int64_t)d.n * d.n;
            if (!D.empty() && (mats + 5 * nn) * 8 > budget) break;
            d.fi = c->h_offs[i];
            d.fj = c->h_offs[j];
            d.oti = oti;
            for (int b = 0; b < 5; ++b) d.buf[b] = mats + b * nn;
            mats += 5 * nn;
            for (int q = 0; q < 2; ++q) {
                d.md[q] = md;
                md += 2 * d.n;
                d.s[q] = sp;
                sp += (int64_t)d.K * d.n;
            }
            d.out = dense;
            dense += (int64_t)M * N;
            nmax = std::max(nmax, d.n);
            Mmax = std::max(Mmax, M);
            D.push_back(d);
            rows.push_back(M);
            cols.push_back(N);
            which.push_back(p);
        }
        const int B = (int)D.size();
        if (B > 0) {
            CU(c->snf_desc.reserve((size_t)B * sizeof(SnfDesc)));
            CU(c->snf_mats.reserve((size_t)mats * 8));
            CU(c->snf_md.reserve((size_t)md * 8));
            CU(c->snf_sidx.reserve((size_t)sp * 4));
            CU(c->snf_sval.reserve((size_t)sp * 8));
            CU(c->dense_in.reserve((size_t)dense * 8));
            CU(cudaMemcpyAsync(c->snf_desc.p, D.data(), (size_t)B * sizeof(SnfDesc), cudaMemcpyHostToDevice, c->stream));
            const SnfDesc* dd = c->snf_desc.as<SnfDesc>();
            double* mt = c->snf_mats.as<double>();
            double* mdp = c->snf_md.as<double>();
            int32_t* si = c->snf_sidx.as<int32_t>();
            double* sv = c->snf_sval.as<double>();
            int rc;
            for (int q = 0; q < 2; ++q) {
                const double* fr = q == 0 ? c->frames.as<double>() : c->frames2.as<double>();
                const int bins = q == 0 ? c->bins : c->bins2;
                if ((rc = launch2d(c, k_snf_dist, nmax, B, dd, fr, mt, bins, (int)m, q, (q == 0 && !no_oti) ? 1 : 0))) return rc;
                if ((rc = launch2d(c, k_snf_meank, nmax, B, dd, mt, mdp, q))) return rc;
                if ((rc = launch2d(c, k_snf_w, nmax, B, dd, mt, (const double*)mdp, q))) return rc;
                if ((rc = launch2d(c, k_snf_s, nmax, B, dd, (const double*)mt, si, sv, q))) return rc;
                if ((rc = launch2d(c, k_snf_regp, nmax, B, dd, mt, q, 2 + q))) return rc;     // P_q from W_q
            }
            // cross diffusion (SimilarityFusion.py:241-271).  cur = current P of each kind, fre = where the next ones go.
            // Iteration 0 reads both OLD matrices; from iteration 1 on `Pts` and `nextPts` are one list in the
            // reference, so kind 1 reads kind 0's matrix of the SAME iteration.
            int cur[2] = {2, 3}, fre[2] = {0, 1};
            for (int it = 0; it < niters; ++it) {
                for (int q = 0; q < 2; ++q) {
                    const int src = (it == 0) ? cur[1 - q] : (q == 0 ? cur[1] : fre[0]);
                    if ((rc = launch2d(c, k_snf_xst, nmax, B, dd, mt, (const int32_t*)si, (const double*)sv, src, q, 4))) return rc;
                    if ((rc = launch2d(c, k_snf_sy, nmax, B, dd, mt, (const int32_t*)si, (const double*)sv, q, 4, fre[q]))) return rc;
                    if ((rc = launch2d(c, k_snf_regp, nmax, B, dd, mt, fre[q], fre[q]))) return rc;
                }
                std::swap(cur[0], fre[0]);
                std::swap(cur[1], fre[1]);
            }
            if ((rc = launch2d(c, k_snf_fused, Mmax, B, dd, (const double*)mt, c->dense_in.as<double>(), cur[0], cur[1]))) return rc;
            if (fused_out) {
                CU(cudaStreamSynchronize(c->stream));
                for (int b = 0; b < B; ++b)
                    CU(cudaMemcpy(fused_out + fused_offsets[which[b]], c->dense_in.as<double>() + D[b].out,
                                  (size_t)D[b].M * D[b].N * 8, cudaMemcpyDeviceToHost));
            }
            std::vector<float> out((size_t)B * 3);
            std::vector<int64_t> co;
            if (crp_out) {
                // the tail writes plot b at co[b]: the caller's offsets of the pairs of this batch
                co.resize(B);
                for (int b = 0; b < B; ++b) co[b] = crp_offsets[which[b]];
            }
            rc = acoss_align_dense_resident(c, ACOSS_F64, rows.data(), cols.data(), B, kappa, flags & ~kFlagNoOti, out.data(),
                                            crp_out, crp_out ? co.data() : nullptr);
            if (rc) return rc;
            for (int b = 0; b < B; ++b) {
                if (qmax_raw) qmax_raw[which[b]] = out[3 * b];
                if (dmax_raw) dmax_raw[which[b]] = out[3 * b + 1];
            }
        }
        p0 = p;
    }
    return ACOSS_OK;
}

}  // extern "C"
