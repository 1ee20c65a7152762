/* CPU oracle, plain C (TEST INFRASTRUCTURE ONLY -- never linked into the product).
 *
 * Restates the transducer loss that /root/reference calls at rnnt/model.py:70
 *     loss = RNNTLoss()(logits, targets.int(), inputs_length.int(), targets_length.int())
 * The arithmetic of that call lives in the third-party, un-vendored, un-pinned
 * package warprnnt_pytorch (HawkAaron/warp-transducer; README.md:4-6), which is not
 * in /root/reference and not installable offline. This file restates the published
 * algorithm (Graves 2012 log-space forward/backward; what warp-transducer's CPU path
 * computes): log-softmax over V, alpha, beta, per-utterance loss, gradient wrt the
 * raw logits. blank index is a parameter; the reference pins blank = 0
 * (model.py:64, decoder.py:9).  PARITY STATUS: parity unpinned against
 * warp-transducer itself (no golden vectors exist in the reference); pinned against
 * torchaudio's independent CPU implementation, the float64 torch DP in
 * rnnt_oracle.py and closed-form cases -- see tests/test_oracle.py.
 *
 * It is also the loss half of bench.py's timed CPU baseline ("port"), so the
 * per-utterance work is OpenMP-parallel over lattice rows.
 *
 * Layouts (all row-major, as the reference passes them): logits (B,T,U1,V) float32,
 * labels (B,U1-1) int32 zero padded, act_lens (B), label_lens (B).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static inline double lae(double a, double b) {
    if (a == -INFINITY) return b;
    if (b == -INFINITY) return a;
    double m = a > b ? a : b;
    return m + log1p(exp(-fabs(a - b)));
}

/* log-sum-exp of one V-vector in float64 */
static double row_lse(const float* x, int V) {
    double m = -INFINITY;
    for (int v = 0; v < V; ++v) if (x[v] > m) m = x[v];
    double s = 0.0;
    for (int v = 0; v < V; ++v) s += exp((double)x[v] - m);
    return m + log(s);
}

/* Returns 0 on success, <0 on invalid argument.
 * loss:  (B) float64, -log P(y|x) per utterance.
 * grad:  nullable; (B,T,U1,V) float32, d loss_b / d logits (unscaled, per utterance),
 *        exactly zero for t >= T_b or u > U_b.
 * c_blank, c_label: nullable (B,T,U1) float64 occupation coefficients (see rnnt_oracle.py). */
int rnnt_oracle_loss(const float* logits, const int32_t* labels, const int32_t* act_lens,
                     const int32_t* label_lens, int B, int T, int U1, int V, int blank,
                     double* loss, float* grad, double* c_blank, double* c_label) {
    if (!logits || !labels || !act_lens || !label_lens || !loss) return -1;
    if (B <= 0 || T <= 0 || U1 <= 0 || V <= 0 || blank < 0 || blank >= V) return -2;
    const int U = U1 - 1;
    for (int b = 0; b < B; ++b) {
        if (act_lens[b] < 1 || act_lens[b] > T) return -3;
        if (label_lens[b] < 0 || label_lens[b] > U) return -4;
        for (int u = 0; u < label_lens[b]; ++u)
            if (labels[(size_t)b * U + u] < 0 || labels[(size_t)b * U + u] >= V) return -5;
    }
    const size_t plane = (size_t)T * U1;
    double* lse = (double*)malloc(sizeof(double) * plane);
    double* lpb = (double*)malloc(sizeof(double) * plane);
    double* lpl = (double*)malloc(sizeof(double) * plane);
    double* al  = (double*)malloc(sizeof(double) * plane);
    double* be  = (double*)malloc(sizeof(double) * plane);
    if (!lse || !lpb || !lpl || !al || !be) { free(lse); free(lpb); free(lpl); free(al); free(be); return -6; }

    for (int b = 0; b < B; ++b) {
        const int Tb = act_lens[b], Ub = label_lens[b];
        const float* xb = logits + (size_t)b * plane * V;
        const int32_t* yb = labels + (size_t)b * U;
        /* log-softmax denominators and the two planes the lattice reads */
        #pragma omp parallel for schedule(static)
        for (int t = 0; t < Tb; ++t) {
            for (int u = 0; u <= Ub; ++u) {
                const float* x = xb + ((size_t)t * U1 + u) * V;
                double z = row_lse(x, V);
                size_t i = (size_t)t * U1 + u;
                lse[i] = z;
                lpb[i] = (double)x[blank] - z;
                lpl[i] = (u < Ub) ? (double)x[yb[u]] - z : -INFINITY;
            }
        }
        /* alpha: alpha[0,0]=0 */
        for (int t = 0; t < Tb; ++t) {
            for (int u = 0; u <= Ub; ++u) {
                size_t i = (size_t)t * U1 + u;
                if (t == 0 && u == 0) { al[i] = 0.0; continue; }
                double a = (t > 0) ? al[i - U1] + lpb[i - U1] : -INFINITY;
                double c = (u > 0) ? al[i - 1] + lpl[i - 1] : -INFINITY;
                al[i] = lae(a, c);
            }
        }
        /* beta: beta[Tb-1,Ub] = lpb[Tb-1,Ub] */
        for (int t = Tb - 1; t >= 0; --t) {
            for (int u = Ub; u >= 0; --u) {
                size_t i = (size_t)t * U1 + u;
                if (t == Tb - 1 && u == Ub) { be[i] = lpb[i]; continue; }
                double a = (t + 1 < Tb) ? be[i + U1] + lpb[i] : -INFINITY;
                double c = (u + 1 <= Ub) ? be[i + 1] + lpl[i] : -INFINITY;
                be[i] = lae(a, c);
            }
        }
        const double ll = be[0];
        loss[b] = -ll;
        if (grad) memset(grad + (size_t)b * plane * V, 0, sizeof(float) * plane * V);
        if (c_blank) memset(c_blank + (size_t)b * plane, 0, sizeof(double) * plane);
        if (c_label) memset(c_label + (size_t)b * plane, 0, sizeof(double) * plane);
        if (grad || c_blank || c_label) {
            #pragma omp parallel for schedule(static)
            for (int t = 0; t < Tb; ++t) {
                for (int u = 0; u <= Ub; ++u) {
                    size_t i = (size_t)t * U1 + u;
                    double cb, cl;
                    if (t + 1 < Tb)      cb = exp(al[i] + lpb[i] + be[i + U1] - ll);
                    else if (u == Ub)    cb = exp(al[i] + lpb[i] - ll);
                    else                 cb = 0.0;
                    cl = (u < Ub) ? exp(al[i] + lpl[i] + be[i + 1] - ll) : 0.0;
                    if (c_blank) c_blank[(size_t)b * plane + i] = cb;
                    if (c_label) c_label[(size_t)b * plane + i] = cl;
                    if (grad) {
                        const float* x = xb + i * V;
                        float* g = grad + ((size_t)b * plane + i) * V;
                        const double c0 = cb + cl, z = lse[i];
                        for (int v = 0; v < V; ++v) g[v] = (float)(exp((double)x[v] - z) * c0);
                        g[blank] = (float)(exp((double)x[blank] - z) * c0 - cb);
                        if (u < Ub) {
                            int y = yb[u];
                            g[y] = (float)(exp((double)x[y] - z) * c0 - cl - (y == blank ? cb : 0.0));
                        }
                    }
                }
            }
        }
    }
    free(lse); free(lpb); free(lpl); free(al); free(be);
    return 0;
}

int rnnt_oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
