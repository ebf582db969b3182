/* oracle/grid_enum.c -- TEST INFRASTRUCTURE ONLY (not product code).
 *
 * Plain-C restatement of the decision-boundary grid of the reference,
 *   robustness/eval_utils.py:31-61  sample_decision_boundary   (the set)
 *   robustness/eval_utils.py:72-89  count_samples_decision_boundary
 * i.e. { z in Z^n_{>=0} : sum z = T, z_0 = max_{i>=1} z_i } as uint8 numerators, rows in the same
 * canonical order as oracle/certify.py:boundary_grid_numerators (shared maximum m ascending, then
 * z_1..z_{n-1} lexicographic).  The pure-Python recursion there is the specification; this file only
 * makes the T = 40 grid (41 320 837 rows) affordable on the CPU.  Pinned in tests/test_oracle_golden.py
 * against the Python recursion (small n, T), against the reference generator's own rows (golden
 * grid_4_8 / grid_10_10) and against count_samples_decision_boundary(10, 40).
 *
 * Built by oracle/build.py:  gcc -O2 -shared -fPIC oracle/grid_enum.c -o oracle/_build/liboracle_grid.so
 */
#include <stddef.h>

typedef struct {
    int n, T, m;
    unsigned char* out; /* NULL: count only */
    long long rows;
    unsigned char cur[16];
} enum_state;

static void rec(enum_state* st, int pos, int remaining, int hit) {
    const int slots = st->n - pos; /* slots still to fill, including this one */
    if (slots == 0) {
        if (remaining == 0 && hit) {
            if (st->out) {
                unsigned char* row = st->out + (size_t)st->rows * st->n;
                for (int j = 0; j < st->n; ++j) row[j] = st->cur[j];
            }
            st->rows++;
        }
        return;
    }
    int lo = remaining - st->m * (slots - 1);
    if (lo < 0) lo = 0;
    const int hi = remaining < st->m ? remaining : st->m;
    for (int z = lo; z <= hi; ++z) {
        st->cur[pos] = (unsigned char)z;
        rec(st, pos + 1, remaining - z, hit || z == st->m);
    }
}

/* Returns the number of rows; writes them to out (rows x n, uint8) unless out is NULL. */
long long oracle_boundary_grid(int n, int T, unsigned char* out) {
    if (n < 2 || n > 16 || T < 1 || T > 255) return -1;
    enum_state st;
    st.n = n; st.T = T; st.out = out; st.rows = 0;
    for (int m = 0; m <= T; ++m) {
        if (T - m < m || (n - 1) * m < T - m) continue;
        st.m = m;
        st.cur[0] = (unsigned char)m;
        rec(&st, 1, T - m, 0);
    }
    return st.rows;
}
