// Host-side check of shifu_b200/csrc/limb.h (built and run by tests/test_limb.py; not part of the product).
// Replays what hist_rows_kernel does to ONE histogram entry -- contributions split into limbs and added to 32-bit words
// that wrap like the hardware's, a scan every 512 rows that normalises the entry when its counter shows >= 512, early
// merges into a 128-bit "global" sum -- against exact 128-bit sums, for adversarial and random contribution streams.
#include <cstdio>
#include <cstdlib>
#include <random>

#include "../shifu_b200/csrc/limb.h"

using namespace sb;
typedef __int128 i128;

struct Entry {
    uint32_t w[LIMB_WORDS] = {0, 0, 0, 0, 0, 0};
    i128 gc = 0, ga = 0, gb = 0;  // what early merges sent to the global accumulators
    i128 ec = 0, ea = 0, eb = 0;  // exact sums
    long merges = 0, normalisations = 0;
    bool ok = true;
    unsigned since = 0;           // contributions since the last normalisation (what bits 22..31 of word 0 must show)

    void add(uint32_t iw, long long a, unsigned long long b) {
        uint32_t d[5];
        limb_split(iw, a, b, d);
        // bounds the analysis in limb.h relies on: checked BEFORE the wrapping add
        const uint64_t c0 = (uint64_t)w[0] + d[0];
        const uint64_t lo_a = (uint64_t)w[1] + d[1], lo_b = (uint64_t)w[3] + d[3], hi_b = (uint64_t)w[4] + d[4];
        const int64_t hi_a = (int64_t)(int32_t)w[2] + (int64_t)(int32_t)d[2];
        if (c0 >> 32 || lo_a >> 32 || lo_b >> 32 || hi_b >> 32 || hi_a > INT32_MAX || hi_a < INT32_MIN) ok = false;
        for (int k = 0; k < 5; ++k) w[k] += d[k];
        ec += iw; ea += a; eb += (i128)b;
        ++since;
        if ((w[0] >> LIMB_BITS) != since) ok = false;
    }
    void scan() {  // what limb_sweep does for this entry
        if (!(w[0] & LIMB_TRIGGER)) return;
        ++normalisations;
        since = 0;
        if (limb_normalise(w, 1)) {
            long long vc, va, vb;
            limb_values(w, 1, vc, va, vb);
            gc += vc; ga += va; gb += (i128)(unsigned long long)vb;
            for (auto& x : w) x = 0;
            ++merges;
        }
    }
    bool final_ok() {
        long long vc, va, vb;
        limb_values(w, 1, vc, va, vb);
        return ok && gc + vc == ec && ga + va == ea && gb + (i128)(unsigned long long)vb == eb;
    }
};

int main() {
    std::mt19937_64 rng(12345);
    const long long AMAX = (1LL << 43) - 1;
    const unsigned long long BMAX = (1ULL << 44) - 1;
    int failures = 0;
    long total_merges = 0, total_norm = 0;
    // stream kinds: 0 all-max positive, 1 all-min negative, 2 alternating extremes, 3 random full range, 4 small values,
    // 5 low-limb-saturating values, 6 random with heavy integer weights
    for (int kind = 0; kind < 7; ++kind) {
        for (int rep = 0; rep < 6; ++rep) {
            Entry e;
            // rows of the block between scans: 512; this entry takes a random share of them (rep 0: every row)
            const long rows = 1 << 18;
            for (long r = 0; r < rows; ++r) {
                const bool hit = rep == 0 || (rng() % (rep == 1 ? 2 : 64)) == 0;
                if (hit) {
                    long long a; unsigned long long b; uint32_t iw = 1;
                    switch (kind) {
                        case 0: a = AMAX; b = BMAX; iw = LIMB_MAX_INT_WEIGHT; break;
                        case 1: a = -AMAX - 1; b = BMAX; break;
                        case 2: a = (r & 1) ? AMAX : -AMAX - 1; b = (r & 1) ? BMAX : 0; break;
                        case 3: a = (long long)(rng() % (2 * (unsigned long long)AMAX + 2)) - AMAX - 1; b = rng() & BMAX; break;
                        case 4: a = (long long)(rng() % 2001) - 1000; b = rng() % 1000; break;
                        case 5: a = (long long)LIMB_MASK; b = LIMB_MASK; break;
                        default: a = (long long)(rng() % (2 * (unsigned long long)AMAX + 2)) - AMAX - 1; b = rng() & BMAX;
                                 iw = (uint32_t)(rng() % (LIMB_MAX_INT_WEIGHT + 1)); break;
                    }
                    e.add(iw, a, b);
                }
                if ((r + 1) % LIMB_SWEEP_ROWS == 0) e.scan();
            }
            if (!e.final_ok()) {
                ++failures;
                printf("FAIL kind %d rep %d (bounds ok: %d)\n", kind, rep, (int)e.ok);
            }
            total_merges += e.merges;
            total_norm += e.normalisations;
        }
    }
    printf("limb_check: %d failures, %ld normalisations, %ld early merges\n", failures, total_norm, total_merges);
    return failures ? 1 : 0;
}
