/*
 * oracle/seqalign_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the three local-alignment recurrences of the reference
 * (benchmarking/SequenceAlignment.c), written from the recurrences in
 * SURVEY.md appendix B.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may call this.
 *
 * Conventions follow the reference so the two can be compared table-for-table:
 *   S  : uint8 binary cross-recurrence plot, row-major
 *   D  : float32 DP table supplied (zeroed) by the caller, left filled on return
 *   ret: maximum table cell as float
 *
 * Every function cites the reference lines it restates.
 */
#include <stddef.h>
#include <stdint.h>

static inline float fmax2(float a, float b) { return b > a ? b : a; }

/* Reference: SequenceAlignment.c:113-143 (qmax_c), gap penalties :104-111
 * (onset == extension == 0.5, so the S[pred] lookups there have no effect).
 * rows = M, cols = N, S[i*N+j]. Cells with i<2 or j<2 are never written. */
float oracle_qmax(const uint8_t *S, float *D, int rows, int cols)
{
    if (rows < 3 || cols < 3) return 0.0f;
    float best = 0.0f;
    for (int i = 2; i < rows; ++i) {
        const float *up1 = D + (size_t)(i - 1) * cols;
        const float *up2 = D + (size_t)(i - 2) * cols;
        float *cur = D + (size_t)i * cols;
        const uint8_t *s = S + (size_t)i * cols;
        for (int j = 2; j < cols; ++j) {
            float p = fmax2(fmax2(up1[j - 1], up2[j - 1]), up1[j - 2]);
            float v = (s[j] == 1) ? p + 1.0f : fmax2(p - 0.5f, 0.0f);
            cur[j] = v;
            best = fmax2(best, v);
        }
    }
    return best;
}

/* Reference: SequenceAlignment.c:147-180 (dmax_c).  Five predecessors, the
 * skipped cells' S values are added as bonuses; penalty 0.5 on a mismatch.
 * Cells with i<3 or j<3 are never written but rows/cols 1..2 ARE read, which
 * is how the reference's table reuse after qmax (Serra09.py:173-175) leaks
 * Qmax values into Dmax.  The caller decides what D holds on entry. */
float oracle_dmax(const uint8_t *S, float *D, int rows, int cols)
{
    if (rows < 4 || cols < 4) return 0.0f;
    float best = 0.0f;
    for (int i = 3; i < rows; ++i) {
        const float *d1 = D + (size_t)(i - 1) * cols;
        const float *d2 = D + (size_t)(i - 2) * cols;
        const float *d3 = D + (size_t)(i - 3) * cols;
        float *cur = D + (size_t)i * cols;
        const uint8_t *s0 = S + (size_t)i * cols;
        const uint8_t *s1 = S + (size_t)(i - 1) * cols;
        const uint8_t *s2 = S + (size_t)(i - 2) * cols;
        for (int j = 3; j < cols; ++j) {
            float c1 = d1[j - 1];
            float c2 = d2[j - 1] + (float)s1[j];
            float c3 = d1[j - 2] + (float)s0[j - 1];
            float c4 = d3[j - 1] + (float)s2[j] + (float)s1[j];
            float c5 = d1[j - 3] + (float)s0[j - 2] + (float)s0[j - 1];
            float p = fmax2(fmax2(fmax2(c1, c2), fmax2(c3, c4)), c5);
            float v = (s0[j] == 1) ? p + 1.0f : fmax2(p - 0.5f, 0.0f);
            cur[j] = v;
            best = fmax2(best, v);
        }
    }
    return best;
}

/* Reference: SequenceAlignment.c:73-99 (swalignimpconstrained), Delta :43-54,
 * Match :56-63.  S is n x m, the table is (n+1) x (m+1); cell (i,j) of the
 * table scores S[i-1][j-1].  The float expression order (pred + match) + gap
 * is kept as the reference writes it. */
static inline float sw_gap(uint8_t before, uint8_t here)
{
    if (here > 0) return 0.0f;
    return before > 0 ? -0.5f : -0.7f;
}

float oracle_sw(const uint8_t *S, float *D, int n, int m)
{
    const int R = n + 1, C = m + 1;
    if (R < 4 || C < 4) return 0.0f;
    float best = 0.0f;
    for (int i = 3; i < R; ++i) {
        for (int j = 3; j < C; ++j) {
            uint8_t here = S[(size_t)(i - 1) * m + (j - 1)];
            float ms = here ? 1.0f : -1.0f;
            float a = D[(size_t)(i - 1) * C + (j - 1)] + ms
                      + sw_gap(S[(size_t)(i - 2) * m + (j - 2)], here);
            float b = D[(size_t)(i - 2) * C + (j - 1)] + ms
                      + sw_gap(S[(size_t)(i - 3) * m + (j - 2)], here);
            float c = D[(size_t)(i - 1) * C + (j - 2)] + ms
                      + sw_gap(S[(size_t)(i - 2) * m + (j - 3)], here);
            float v = fmax2(fmax2(a, b), fmax2(c, 0.0f));
            D[(size_t)i * C + j] = v;
            best = fmax2(best, v);
        }
    }
    return best;
}

/* Serra09.py:173-175 in one call: zeroed table -> qmax -> dmax on the SAME
 * table (the "stale boundary" quirk). */
void oracle_qmax_then_dmax(const uint8_t *S, float *D, int rows, int cols,
                           float *qmax_out, float *dmax_out)
{
    for (size_t t = 0; t < (size_t)rows * cols; ++t) D[t] = 0.0f;
    *qmax_out = oracle_qmax(S, D, rows, cols);
    *dmax_out = oracle_dmax(S, D, rows, cols);
}
