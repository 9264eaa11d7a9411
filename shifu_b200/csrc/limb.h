// limb.h -- the integer arithmetic of the limb accumulation used by hist_rows_kernel (hist.cu), as plain inline
// functions so that the same code runs in the kernel and in a host-side exhaustive check (tests/limb_check.cpp).
//
// A returning shared-memory atomic costs the LSU about 1.85x a fire-and-forget one (tools/ubench_smem.cu), and a
// 64-bit sum built from 32-bit atomics needs one per statistic to learn the carry.  Instead a contribution of up to
// 44 bits is cut into two 22-bit limbs and each limb is added to its own 32-bit word with RED (no return, no carry):
// 10 bits of headroom, i.e. up to 1023 contributions per word before it must be normalised.  One (slot, bin, feature)
// entry is six words:
//     word 0   bits 0..21  sum of the integer sig*weight products      bits 22..31  contributions since normalised
//     word 1/2 sums of the low / high limbs of y*sig*w    (high limb signed; the statistic is scaled to |q| < 2^43)
//     word 3/4 sums of the low / high limbs of y*y*sig*w  (unsigned, q < 2^44)
//     word 5   bits 0..15 signed / bits 16..31 unsigned: what normalisation carried out of words 2 / 4
// Every 512 rows the consumer warps scan the count words; an entry with bit 31 set (>= 512 contributions) is
// normalised: low-limb carries move to the high word, high-word carries to word 5, the contribution counter restarts.
// Between two scans an entry gains at most 512 contributions, so no counter passes 1023 and no word wraps.  When a
// field of word 5 or the weight sum is half full the entry is merged into the global accumulators early.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define SB_HD __host__ __device__ __forceinline__
#else
#define SB_HD inline
#endif

namespace sb {

constexpr int LIMB_BITS = 22;
constexpr uint32_t LIMB_MASK = (1u << LIMB_BITS) - 1u;
constexpr uint32_t LIMB_MARK = 1u << LIMB_BITS;   // one contribution, in the count word
constexpr uint32_t LIMB_TRIGGER = 1u << 31;       // 512 contributions
constexpr int LIMB_SWEEP_ROWS = 512;              // rows between two scans
constexpr int LIMB_WORDS = 6;
constexpr int LIMB_HI_KEEP = 21;                  // bits of the signed high word kept at normalisation
constexpr int LIMB_MAX_INT_WEIGHT = 1024;         // check_int_weights_kernel: sig*w is an integer in [0, 1024]
static_assert((uint64_t)LIMB_SWEEP_ROWS << LIMB_BITS == LIMB_TRIGGER, "the trigger is bit 31 of the count word");

// the five addends of one contribution: integer weight iw, sum operand a (|a| < 2^43), square operand b (0 <= b < 2^44)
SB_HD void limb_split(uint32_t iw, long long a, unsigned long long b, uint32_t out[5]) {
    out[0] = iw + LIMB_MARK;
    out[1] = (uint32_t)a & LIMB_MASK;
    out[2] = (uint32_t)(a >> LIMB_BITS);  // arithmetic shift: signed high limb in [-2^21, 2^21)
    out[3] = (uint32_t)b & LIMB_MASK;
    out[4] = (uint32_t)(b >> LIMB_BITS);
}

// value of an entry whose words are `stride` apart: weight sum, sum statistic, square statistic (fixed point)
SB_HD void limb_values(const uint32_t* p, int stride, long long& vc, long long& va, long long& vb) {
    const uint32_t ov = p[5 * stride];
    const long long ova = (long long)(int16_t)(ov & 0xffffu);
    const unsigned long long ovb = ov >> 16;
    vc = (long long)(p[0] & LIMB_MASK);
    // (ova * 2^21 + high) * 2^22 + low, as multiplications: the partial sums may be negative
    va = (ova * (1LL << LIMB_HI_KEEP) + (long long)(int32_t)p[2 * stride]) * (1LL << LIMB_BITS) + (long long)p[stride];
    vb = (long long)((((ovb << LIMB_BITS) + (unsigned long long)p[4 * stride]) << LIMB_BITS) + (unsigned long long)p[3 * stride]);
}

// Normalises one entry in place (the caller holds it exclusively).  Returns true when the entry should be merged into
// the global accumulators and cleared now (a carry field or the weight sum is half full).
SB_HD bool limb_normalise(uint32_t* p, int stride) {
    const uint32_t siw = p[0] & LIMB_MASK;
    uint32_t alo = p[stride], ahi = p[2 * stride], blo = p[3 * stride], bhi = p[4 * stride];
    const uint32_t ov = p[5 * stride];
    ahi += alo >> LIMB_BITS;
    alo &= LIMB_MASK;
    const int32_t ca = (int32_t)ahi >> LIMB_HI_KEEP;  // floor: the kept part is in [0, 2^21)
    ahi -= (uint32_t)ca << LIMB_HI_KEEP;
    const int32_t ova = (int32_t)(int16_t)(ov & 0xffffu) + ca;
    bhi += blo >> LIMB_BITS;
    blo &= LIMB_MASK;
    const uint32_t ovb = (ov >> 16) + (bhi >> LIMB_BITS);
    bhi &= LIMB_MASK;
    p[0] = siw;
    p[stride] = alo;
    p[2 * stride] = ahi;
    p[3 * stride] = blo;
    p[4 * stride] = bhi;
    p[5 * stride] = ((uint32_t)ova & 0xffffu) | (ovb << 16);
    return ova >= 16384 || ova <= -16384 || ovb >= 32768u || siw >= (1u << 21);
}

}  // namespace sb
