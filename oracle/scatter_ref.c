/*
 * ORACLE (test infrastructure only; parity unpinned — see oracle/ref_cpu.py).
 *
 * Plain-C restatement of the sequential CPU loops of torch-scatter==2.0.4
 * (csrc/cpu/scatter_cpu.cpp; pinned at /root/reference/requirement.txt:3, not
 * vendored in the reference tree) as described in SURVEY.md Appendix A.4-A.6,
 * and of the gather -> message -> scatter pipeline of
 * MessagePassing.propagate for GCN-style messages (A.1/A.9; reference call
 * sites examples/gcn.py:25,27).  Single-threaded, edge order = input order.
 * Used by tests to cross-check oracle/ref_cpu.py and by bench.py's
 * cpu_baseline as a scalar port.
 */
#include <float.h>
#include <stdint.h>
#include <string.h>

/* A.4: out[index[e], k] += src[e, k]  (out zero-filled by the caller) */
void ref_scatter_add(const float* src, const int64_t* index, int64_t E, int64_t K, float* out) {
  for (int64_t e = 0; e < E; ++e) {
    float* o = out + index[e] * K;
    const float* s = src + e * K;
    for (int64_t k = 0; k < K; ++k) o[k] += s[k];
  }
}

/* A.5: mean = sum / max(count, 1) */
void ref_scatter_mean(const float* src, const int64_t* index, int64_t E, int64_t K, int64_t N,
                      float* out, float* count /* [N] scratch */) {
  memset(out, 0, sizeof(float) * (size_t)(N * K));
  memset(count, 0, sizeof(float) * (size_t)N);
  ref_scatter_add(src, index, E, K, out);
  for (int64_t e = 0; e < E; ++e) count[index[e]] += 1.0f;
  for (int64_t n = 0; n < N; ++n) {
    float c = count[n] < 1.0f ? 1.0f : count[n];
    for (int64_t k = 0; k < K; ++k) out[n * K + k] /= c;
  }
}

/* A.6: strict '>' => first maximal edge wins; empty rows: value 0, arg = E */
void ref_scatter_max(const float* src, const int64_t* index, int64_t E, int64_t K, int64_t N,
                     float* out, int64_t* arg) {
  for (int64_t i = 0; i < N * K; ++i) {
    out[i] = -FLT_MAX;
    arg[i] = E;
  }
  for (int64_t e = 0; e < E; ++e) {
    int64_t n = index[e];
    for (int64_t k = 0; k < K; ++k) {
      float v = src[e * K + k];
      if (v > out[n * K + k]) {
        out[n * K + k] = v;
        arg[n * K + k] = e;
      }
    }
  }
  for (int64_t i = 0; i < N * K; ++i)
    if (arg[i] == E) out[i] = 0.0f;
}

/* A.1 + A.9: out[col[e], :] += norm[e] * x[row[e], :], materialising the
 * message row like the reference does (index_select, mul, scatter_add). */
void ref_propagate_gcn(const float* x, const int64_t* row, const int64_t* col, const float* norm,
                       int64_t E, int64_t K, float* msg /* [K] scratch */, float* out) {
  for (int64_t e = 0; e < E; ++e) {
    const float* xr = x + row[e] * K;
    float w = norm ? norm[e] : 1.0f;
    for (int64_t k = 0; k < K; ++k) msg[k] = w * xr[k];
    float* o = out + col[e] * K;
    for (int64_t k = 0; k < K; ++k) o[k] += msg[k];
  }
}
