/*
 * gms_oracle.c -- CPU restatement of the GMS (grid-based motion statistics) match filter as the reference calls it.
 *
 * TEST INFRASTRUCTURE ONLY (checker for tests/ and bench.py; nothing under matchinglib_poselib_b200/ may link it).
 *
 * Reference (paths relative to /root/reference/matchinglib_poselib/source/matchinglib):
 *   source/gms.cpp:54-60                       filterMatchesGMS -> MatchGMS(...).getInlierMask(mask, useScale, useRotation)
 *   source/correspondences.cpp:378-396         call site: imgSi for both images, useScale = useRotation = false
 *   thirdparty/gms-1.0/src/MatchGMS.cpp        the algorithm:
 *     :55-75   constructor (20 x 20 left grid)        :77-101  getInlierMask (best of scales x rotations, strict '>')
 *     :103-116 normalizeKeypoints (float, * 1/size)   :139-171 run (4 shifted grid types, inliers OR-ed)
 *     :173-201 assignMatchPairs (right cell fixed by grid type 1; a match that falls off a grid stays invalid)
 *     :203-248 verifyCellPairs (first arg-max per left cell; 3 x 3 neighbourhood score against 6 sqrt(mean cell population))
 *     :250-259 setScale        :261-281 getN9        :283-318 getGridIndexLeft / getGridIndexRight
 * Deviation (documented): a right-grid index outside [0, cells) -- a keypoint on or beyond the image border -- indexes past
 * the reference's matrix (undefined behaviour there); here such a match is treated as invalid (-1, -1).
 *
 * Parity pin: oracle/_ref/libgms_ref.so is the reference's own MatchGMS.cpp compiled unmodified (oracle/Makefile ref_gms);
 * tests/test_oracle_gms.py checks this file against it on seeded inputs and against tests/golden/gms_golden.npz.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static const int ROT[8][9] = {
    {1,2,3,4,5,6,7,8,9}, {4,1,2,7,5,3,8,9,6}, {7,4,1,8,5,2,9,6,3}, {8,7,4,9,5,1,6,3,2},
    {9,8,7,6,5,4,3,2,1}, {6,9,8,3,5,7,2,1,4}, {3,6,9,2,5,8,1,4,7}, {2,3,6,1,5,9,4,7,8}};

static void n9(int cell, int gw, int gh, int* out)
{
    for (int k = 0; k < 9; ++k) out[k] = -1;
    const int ix = cell % gw, iy = cell / gw;
    for (int yi = -1; yi <= 1; ++yi)
        for (int xi = -1; xi <= 1; ++xi) {
            const int xx = ix + xi, yy = iy + yi;
            if (xx < 0 || xx >= gw || yy < 0 || yy >= gh) continue;
            out[xi + 4 + yi * 3] = xx + yy * gw;
        }
}

static int grid_left(float x, float y, int type, int gw, int gh)
{
    int ix, iy;
    if (type == 1)      { ix = (int)floorf(x * gw);        iy = (int)floorf(y * gh); }
    else if (type == 2) { ix = (int)floorf(x * gw + 0.5f); iy = (int)floorf(y * gh); }
    else if (type == 3) { ix = (int)floorf(x * gw);        iy = (int)floorf(y * gh + 0.5f); }
    else                { ix = (int)floorf(x * gw + 0.5f); iy = (int)floorf(y * gh + 0.5f); }
    if (ix >= gw || iy >= gh) return -1;
    return ix + iy * gw;
}

/* kp*: n x 2 float pixel coordinates; matches: n_matches x 2 (queryIdx, trainIdx); mask out.  Returns the inlier count. */
int orc_gms_filter(const float* kp1, int n1, int w1, int h1, const float* kp2, int n2, int w2, int h2,
                   const int32_t* matches, int nm, int use_scale, int use_rotation, uint8_t* mask_out)
{
    (void)n1; (void)n2;
    static const double scale_ratio[5] = {1.0, 0.5, 0.70710678118654757, 1.4142135623730951, 2.0};
    const int gwl = 20, ghl = 20, cells_l = gwl * ghl;
    const float w1i = 1.0f / w1, h1i = 1.0f / h1, w2i = 1.0f / w2, h2i = 1.0f / h2;
    const int n_scales = use_scale ? 5 : 1, n_rot = use_rotation ? 8 : 1;
    int* pl = (int*)malloc(sizeof(int) * (nm > 0 ? nm : 1)); int* pr = (int*)malloc(sizeof(int) * (nm > 0 ? nm : 1));
    uint8_t* cur = (uint8_t*)malloc(nm > 0 ? nm : 1);
    int* cell_pairs = (int*)malloc(sizeof(int) * cells_l); int* npts = (int*)malloc(sizeof(int) * cells_l);
    int best = 0;
    memset(mask_out, 0, nm);
    for (int sl = 0; sl < n_scales; ++sl) {
        const int gwr = (int)(gwl * scale_ratio[sl]), ghr = (int)(ghl * scale_ratio[sl]), cells_r = gwr * ghr;
        int* stat = (int*)malloc(sizeof(int) * cells_l * cells_r);
        for (int rt = 0; rt < n_rot; ++rt) {
            memset(cur, 0, nm);
            for (int i = 0; i < nm; ++i) { pl[i] = 0; pr[i] = 0; }
            for (int type = 1; type <= 4; ++type) {
                memset(stat, 0, sizeof(int) * cells_l * cells_r);
                for (int c = 0; c < cells_l; ++c) { cell_pairs[c] = -1; npts[c] = 0; }
                for (int i = 0; i < nm; ++i) {                                   /* assignMatchPairs */
                    const float lx = kp1[2 * matches[2 * i]] * w1i, ly = kp1[2 * matches[2 * i] + 1] * h1i;
                    const float rx = kp2[2 * matches[2 * i + 1]] * w2i, ry = kp2[2 * matches[2 * i + 1] + 1] * h2i;
                    int l = pl[i] = grid_left(lx, ly, type, gwl, ghl), r;
                    if (type == 1) {
                        r = (int)floorf(rx * gwr) + (int)floorf(ry * ghr) * gwr;
                        if (r >= cells_r) r = -1;                                /* see "Deviation" above */
                        pr[i] = r;
                    } else r = pr[i];
                    if (l < 0 || r < 0) { pl[i] = pr[i] = -1; continue; }
                    stat[l * cells_r + r]++;
                    npts[l]++;
                }
                for (int c = 0; c < cells_l; ++c) {                             /* verifyCellPairs */
                    long rowsum = 0; int mx = 0;
                    for (int j = 0; j < cells_r; ++j) rowsum += stat[c * cells_r + j];
                    if (rowsum == 0) { cell_pairs[c] = -1; continue; }
                    for (int j = 0; j < cells_r; ++j) if (stat[c * cells_r + j] > mx) { cell_pairs[c] = j; mx = stat[c * cells_r + j]; }
                    int ln[9], rn[9];
                    n9(c, gwl, ghl, ln); n9(cell_pairs[c], gwr, ghr, rn);
                    int score = 0, npair = 0; double th = 0;
                    for (int j = 0; j < 9; ++j) {
                        const int ll = ln[j], rr = rn[ROT[rt][j] - 1];
                        if (ll == -1 || rr == -1) continue;
                        score += stat[ll * cells_r + rr]; th += npts[ll]; ++npair;
                    }
                    th = 6 * sqrt(th / npair);
                    if (score < th) cell_pairs[c] = -2;
                }
                for (int i = 0; i < nm; ++i) {                                   /* mark inliers */
                    if (pl[i] == -1 && pr[i] == -1) continue;
                    if (cell_pairs[pl[i]] == pr[i]) cur[i] = 1;
                }
            }
            int cnt = 0;
            for (int i = 0; i < nm; ++i) cnt += cur[i];
            if (cnt > best) { best = cnt; memcpy(mask_out, cur, nm); }
        }
        free(stat);
    }
    free(pl); free(pr); free(cur); free(cell_pairs); free(npts);
    return best;
}
