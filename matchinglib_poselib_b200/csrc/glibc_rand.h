// glibc_rand.h -- glibc's srand() / rand() restated (host only), shared by usac.cu and vfc.cu: the reference draws its USAC samples,
// evaluation-pool shuffle and SparseVFC control points from rand(), so the stream of a seed here IS the reference's.
#pragma once
#include <stdint.h>

// glibc's srand() / rand() (TYPE_3 additive feedback generator, x^31 + x^3 + 1): the reference seeds it in
// estimateEssentialMatUsac (usac_estimations.cpp:328) and draws every sample and the evaluation-pool shuffle from it;
// VFC::selectSubset (matchinglib/source/vfc.cpp:136) draws from the process' current state (srand(1) in a fresh process).
struct GlibcRand {
    uint32_t st[31]; int f = 3, r = 0;
    void seed(unsigned s)
    {
        if (s == 0) s = 1;
        int32_t word = (int32_t)s;
        st[0] = (uint32_t)word;
        for (int i = 1; i < 31; ++i) {
            const long hi = word / 127773, lo = word % 127773;
            long w = 16807 * lo - 2836 * hi;
            if (w < 0) w += 2147483647;
            word = (int32_t)w; st[i] = (uint32_t)word;
        }
        f = 3; r = 0;
        for (int i = 0; i < 310; ++i) next();
    }
    int next()
    {
        st[f] += st[r];
        const int res = (int)(st[f] >> 1);
        if (++f >= 31) f = 0;
        if (++r >= 31) r = 0;
        return res;
    }
};

