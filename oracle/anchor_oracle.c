/*
 * anchor_oracle.c -- CPU restatement of the reference's empty-anchor filter (TEST INFRASTRUCTURE ONLY; nothing
 * under mv3d_b200/ may link or call this).
 *
 * Follows src/net/utility/remove_empty_box_kernel.cu:3-42 (`integration`, `remove_empty`) as launched by
 * src/net/utility/remove_empty_box.py:25-47: grid = (anchors, 50), block = (channels): block (a, line) adds the
 * float32 sum of one anchor row -- view[x1+line][y1 .. y2)[channel], accumulated left to right from 0 -- into
 * inds[a][channel]; the host then keeps the anchors whose np.sum(inds, axis=1) exceeds `limit`.
 *
 * The reference's `+=` into inds[] from up to 50 blocks is not atomic (remove_empty_box_kernel.cu:41), so its own
 * result is not deterministic. This restatement is the race-free schedule "blocks retire in ascending blockIdx.y":
 * one of the outcomes the reference can produce, and the only one it produces when nothing is lost.
 *
 * Parity status: the row integration, the clamping and the 50-line cap are pinned by EXECUTING the reference kernel
 * (compiled from where it lies by oracle/Makefile into oracle/_ref/remove_empty_box_kernel.cubin, launched on the GPU
 * box by tests/test_gpu_anchors.py) on anchors one line wide, where the race cannot occur; wider anchors are checked
 * on integer-valued views, where every schedule without a lost update gives the same sum.
 *
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC anchor_oracle.c -o _build/libanchor_oracle.so
 */
#include <stdint.h>

#define MAX_LINES 50 /* grid=(anchors, 50, 1), remove_empty_box.py:43 */

static int clampi(int v, int n) /* remove_empty_box_kernel.cu:22-29 */
{
    if (v < 0) v = 0;
    if (v >= n) v = n - 1;
    return v;
}

/* inds[n][c] for one view[xn][yn][c]; anchors[n][4] = (y1, x1, y2, x2), x indexing view dimension 0. */
void anchor_channel_sums(const float *view, int xn, int yn, int c, const int32_t *anchors, long n, float *inds)
{
    for (long a = 0; a < n; a++) {
        const int32_t *b = anchors + 4 * a;
        const int y1 = clampi(b[0], yn), x1 = clampi(b[1], xn), y2 = clampi(b[2], yn), x2 = clampi(b[3], xn);
        const int w = x2 - x1, l = y2 - y1;
        float *out = inds + a * c;
        for (int ch = 0; ch < c; ch++) out[ch] = 0.0f;
        for (int line = 0; line < w && line < MAX_LINES; line++) {
            const float *row = view + ((long)(x1 + line) * yn + y1) * c;
            for (int ch = 0; ch < c; ch++) {
                float sum = 0.0f; /* `integration`, :3-10 */
                for (int i = 0; i < l; i++) sum += row[(long)i * c + ch];
                out[ch] += sum; /* :41 */
            }
        }
    }
}
