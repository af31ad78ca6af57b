// k_paths_packed.cuh -- the path-aggregation kernels in packed int16x2 arithmetic (included by k_sgbm.cu).
//
// sm_100a executes VIADD.16x2, VIMNMX(3).S16x2 and VIADDMNMX.S16x2/U16x2 natively, so two disparities share one
// 32-bit lane operation and costs never need unpacking.  The packed ops wrap instead of saturating, therefore every
// int16 saturation of OpenCV's lanes is restated so that no intermediate can leave the int16 range:
//   sat(Lp[d+-1] + P1s)       -> min(Lp[d-1], Lp[d+1], 32767 - P1s) + P1s          (P1s >= 0; mirrored when P1s < 0)
//   sat(m - delta), m <= delta -> max(m, max(delta - 32768, -32768)) - delta
//   sat(t + C), t <= 0         -> t + C when C >= 0 (frame flag bit 0 clear), else max(C, -32768 - t) + t
//   sat(S + L)                 -> min(S + L, 32767) in unsigned lanes while all L >= 0 so far (flag bit 1 clear),
//                                 else the generic saturating add
// The two flag bits are per frame, live in global memory and are sticky: k_vsum raises bit 0 when it writes a
// negative C, a path kernel raises bit 1 the moment one of its columns produces a negative path cost (and switches
// itself to the exact form first).  In the reference's parameter regimes neither ever fires; the adversarial
// `satdrift` fixture exercises both.  Results are bit-identical to the int32 kernels (k_path_h / k_path_v).
#pragma once

__device__ __forceinline__ unsigned pk2(int v) { return ((unsigned)v & 0xffffu) * 0x10001u; }

struct PathK {                 // per-launch constants of the packed update
    unsigned P1p;              // P1s in both halves
    unsigned Qp;               // 32767 - P1s (P1s >= 0) or -32768 - P1s (P1s < 0)
    unsigned Z;                // 0, kept in a register (a literal zero operand of min.s16x2 costs a PRMT per use)
};

// EXACT = false assumes C >= 0 (frame flag bit 0 clear); EXACT = true takes the general form of t + C.
template <bool EXACT>
__device__ __forceinline__ unsigned path_cost2(unsigned lm1, unsigned l0, unsigned lp1, unsigned deltap, unsigned lop, unsigned ndeltap,
                                               const PathK& k, unsigned c)
{
    const unsigned n = __vimin3_s16x2(lm1, lp1, k.Qp);               // min(Lp[d-1], Lp[d+1], 32767 - P1s)
    const unsigned t = __viaddmin_s16x2(n, k.P1p, l0);               // min(n + P1s, Lp[d])
    const unsigned m = __vmaxs2(__vmins2(t, deltap), lop);           // min(., delta), kept >= delta - 32768
    const unsigned t2 = __vadd2(m, ndeltap);                         // m - delta in [-32768, 0]
    if (!EXACT) return __vadd2(t2, c);
    const unsigned w = __vadd2(~t2, 0x80018001u);                    // -32768 - t2
    return __vadd2(__vmaxs2(c, w), t2);                              // max(t2 + C, -32768)
}

template <int NWD>
__device__ __forceinline__ void ldw(const unsigned* p, unsigned (&v)[NWD])
{
    if (NWD == 2) { const uint2 a = *(const uint2*)p; v[0] = a.x; v[1] = a.y; }
    else if (NWD == 4) { const uint4 a = *(const uint4*)p; v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; }
    else {
#pragma unroll
        for (int i = 0; i < NWD; i++) v[i] = p[i];
    }
}
template <int NWD>
__device__ __forceinline__ void stw(unsigned* p, const unsigned (&v)[NWD])
{
    if (NWD == 2) *(uint2*)p = make_uint2(v[0], v[1]);
    else if (NWD == 4) *(uint4*)p = make_uint4(v[0], v[1], v[2], v[3]);
    else {
#pragma unroll
        for (int i = 0; i < NWD; i++) p[i] = v[i];
    }
}

// ---- asynchronous prefetch ring (cp.async, Ampere+ LDGSTS): the warp copies the C and S words of a column several
// columns ahead into shared memory, so the long global-load latency is off the per-column critical path.
// cpa_words: a lane copies its own NWD words (a lane then only reads back bytes it copied itself, cp.async.wait_group is
// all the synchronisation needed).  cpa_16: 16-byte chunks (the k_path_h2 ring: the first Dp/8 lanes cover the column of
// C, the next Dp/8 the column of S, one instruction for both arrays when Dp <= 128; needs a __syncwarp after the wait).
template <int NWD>
__device__ __forceinline__ void cpa_words(unsigned* sdst, const unsigned* gsrc)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(sdst);
    if (NWD == 2) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gsrc) : "memory");
    else if (NWD == 4) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
    else {
#pragma unroll
        for (int i = 0; i < NWD; i++) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 4 * i), "l"(gsrc + i) : "memory");
    }
}
__device__ __forceinline__ void cpa_16(unsigned sdst, const void* gsrc)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sdst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cpa_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cpa_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
#define CVR_PF_DEPTH 4          // columns in flight per warp

// One column update for one direction: Lp (previous costs of the source pixel, packed) -> L (this pixel); returns
// the warp-wide minimum of L.  EXACT = true: general form (any sign of C, Lp, delta).  EXACT = false: the caller
// knows C >= 0 for this frame; if in addition the source costs are non-negative (minPrev >= 0) and delta >= 0 --
// which is every column outside the wrap/saturation regimes -- the update shrinks to
//     t = min(min(Lp[d-1], Lp[d+1], 32767 - P1s) + P1s, Lp[d]);  L = min(t - delta, 0) + C
// (t in [0, 32767] and delta in [0, 32767], so t - delta cannot wrap); otherwise it takes the clamped form.
template <int NWD, bool EXACT, bool HASPAD>
__device__ __forceinline__ int column_update(const unsigned (&Lp)[NWD], int minPrev, const unsigned (&c)[NWD], unsigned (&L)[NWD], int lane, int D,
                                             int P2, const PathK& k)
{
    const int delta = (int)(short)(P2 + minPrev);
    unsigned left = __shfl_up_sync(0xffffffffu, Lp[NWD - 1], 1);
    unsigned right = __shfl_down_sync(0xffffffffu, Lp[0], 1);
    if (lane == 0) left = 0x7fff7fffu;                 // Lp[-1] = MAX_COST
    if (lane == 31) right = 0x7fff7fffu;               // Lp[D]  = MAX_COST (padding lanes hold MAX_COST too)
    const unsigned ndeltap = pk2(-delta);
    unsigned mnp = 0x7fff7fffu;
    if (!EXACT && (minPrev | delta) >= 0) {
#pragma unroll
        for (int j = 0; j < NWD; j++) {
            const unsigned lm1 = __byte_perm(j == 0 ? left : Lp[j - 1], Lp[j], 0x5432);           // (Lp[2j-1], Lp[2j])
            const unsigned lp1 = __byte_perm(Lp[j], j == NWD - 1 ? right : Lp[j + 1], 0x5432);     // (Lp[2j+1], Lp[2j+2])
            const unsigned n = __vimin3_s16x2(lm1, lp1, k.Qp);
            const unsigned t = __viaddmin_s16x2(n, k.P1p, Lp[j]);
            unsigned v = __vadd2(__viaddmin_s16x2(t, ndeltap, 0u), c[j]);
            if (HASPAD && (lane * NWD + j) * 2 >= D) v = 0x7fff7fffu;
            L[j] = v;
            mnp = __vmins2(mnp, v);
        }
    } else {
        const unsigned deltap = pk2(delta), lop = pk2(max(delta - 32768, -32768));
#pragma unroll
        for (int j = 0; j < NWD; j++) {
            const unsigned lm1 = __byte_perm(j == 0 ? left : Lp[j - 1], Lp[j], 0x5432);
            const unsigned lp1 = __byte_perm(Lp[j], j == NWD - 1 ? right : Lp[j + 1], 0x5432);
            unsigned v;
            if (EXACT) v = path_cost2<true>(lm1, Lp[j], lp1, deltap, lop, ndeltap, k, c[j]);
            else v = path_cost2<false>(lm1, Lp[j], lp1, deltap, lop, ndeltap, k, c[j]);
            if (HASPAD && (lane * NWD + j) * 2 >= D) v = 0x7fff7fffu;
            L[j] = v;
            mnp = __vmins2(mnp, v);
        }
    }
    const int m = min((int)(short)(mnp & 0xffffu), ((int)mnp) >> 16);
    return __reduce_min_sync(0xffffffffu, m);
}

template <int NWD>
__device__ __forceinline__ void acc_fast(unsigned (&s)[NWD], const unsigned (&L)[NWD])        // S, L >= 0: min(S + L, 32767) in unsigned lanes
{
#pragma unroll
    for (int j = 0; j < NWD; j++) s[j] = __viaddmin_u16x2(s[j], L[j], 0x7fff7fffu);
}
template <int NWD>
__device__ __forceinline__ void acc_exact(unsigned (&s)[NWD], const unsigned (&L)[NWD])       // generic saturating int16 add
{
#pragma unroll
    for (int j = 0; j < NWD; j++) s[j] = __vaddss2(s[j], L[j]);
}

// Fast column update (preconditions checked by the caller, all warp-uniform: C >= 0 in this frame, source costs
// non-negative, 0 <= delta <= 32767): returns this lane's minimum of L (not yet reduced over the warp).
// NN ("no neighbours"): when 0 <= P2s <= P1s (P2s = (short)P2 -- the reference's own default, block_size 27, lands
// here because 96*27^2 wraps to 4448 while P1 stays 17496) and delta did not wrap (delta == P2s + min), the two
// neighbour terms Lp[d+-1] + P1s >= min + P1s >= min + P2s = delta can never win the four-way minimum, so the update
// is exactly  L = min(Lp[d] - delta, 0) + C  -- no shuffles, no permutes.
template <int NWD, bool HASPAD, bool NN>
__device__ __forceinline__ unsigned col_fast_pk(const unsigned (&Lp)[NWD], unsigned ndeltap, const unsigned (&c)[NWD], unsigned (&L)[NWD], int lane, int D,
                                        const PathK& k)
{
    unsigned mnp = 0x7fff7fffu;
    if (NN) {
#pragma unroll
        for (int j = 0; j < NWD; j++) {
            unsigned v = __vadd2(__viaddmin_s16x2(Lp[j], ndeltap, k.Z), c[j]);
            if (HASPAD && (lane * NWD + j) * 2 >= D) v = 0x7fff7fffu;
            L[j] = v;
            mnp = __vmins2(mnp, v);
        }
    } else {
        unsigned left = __shfl_up_sync(0xffffffffu, Lp[NWD - 1], 1);
        unsigned right = __shfl_down_sync(0xffffffffu, Lp[0], 1);
        if (lane == 0) left = 0x7fff7fffu;
        if (lane == 31) right = 0x7fff7fffu;
#pragma unroll
        for (int j = 0; j < NWD; j++) {
            const unsigned lm1 = __byte_perm(j == 0 ? left : Lp[j - 1], Lp[j], 0x5432);
            const unsigned lp1 = __byte_perm(Lp[j], j == NWD - 1 ? right : Lp[j + 1], 0x5432);
            const unsigned n = __vimin3_s16x2(lm1, lp1, k.Qp);
            const unsigned t = __viaddmin_s16x2(n, k.P1p, Lp[j]);
            unsigned v = __vadd2(__viaddmin_s16x2(t, ndeltap, k.Z), c[j]);
            if (HASPAD && (lane * NWD + j) * 2 >= D) v = 0x7fff7fffu;
            L[j] = v;
            mnp = __vmins2(mnp, v);
        }
    }
    return mnp;                                                       // the lane's two candidate minima, packed
}
template <int NWD, bool HASPAD, bool NN>
__device__ __forceinline__ int col_fast(const unsigned (&Lp)[NWD], unsigned ndeltap, const unsigned (&c)[NWD], unsigned (&L)[NWD], int lane, int D,
                                        const PathK& k)
{
    const unsigned mnp = col_fast_pk<NWD, HASPAD, NN>(Lp, ndeltap, c, L, lane, D, k);
    return min((int)(short)(mnp & 0xffffu), ((int)mnp) >> 16);
}
// General column update for a given delta (any sign of Lp, delta; EXACT also any sign of C).
template <int NWD, bool EXACT, bool HASPAD>
__device__ __forceinline__ int col_general(const unsigned (&Lp)[NWD], int delta, const unsigned (&c)[NWD], unsigned (&L)[NWD], int lane, int D, const PathK& k)
{
    unsigned left = __shfl_up_sync(0xffffffffu, Lp[NWD - 1], 1);
    unsigned right = __shfl_down_sync(0xffffffffu, Lp[0], 1);
    if (lane == 0) left = 0x7fff7fffu;
    if (lane == 31) right = 0x7fff7fffu;
    const unsigned deltap = pk2(delta), lop = pk2(max(delta - 32768, -32768)), ndeltap = pk2(-delta);
    unsigned mnp = 0x7fff7fffu;
#pragma unroll
    for (int j = 0; j < NWD; j++) {
        const unsigned lm1 = __byte_perm(j == 0 ? left : Lp[j - 1], Lp[j], 0x5432);
        const unsigned lp1 = __byte_perm(Lp[j], j == NWD - 1 ? right : Lp[j + 1], 0x5432);
        unsigned v = path_cost2<EXACT>(lm1, Lp[j], lp1, deltap, lop, ndeltap, k, c[j]);
        if (HASPAD && (lane * NWD + j) * 2 >= D) v = 0x7fff7fffu;
        L[j] = v;
        mnp = __vmins2(mnp, v);
    }
    return min((int)(short)(mnp & 0xffffu), ((int)mnp) >> 16);
}
// What the next pixel on a path needs to know about this pixel's minimum, precomputed by the producer:
//   A = -delta in both halves, B = (delta & 0xffff) | 0x80000000 if the fast form is not allowed (min < 0 or delta < 0)
template <bool NN>
__device__ __forceinline__ void path_info(int mn, int P2, unsigned& A, unsigned& B)
{
    const int delta = (int)(short)(P2 + mn);
    A = pk2(-delta);
    int bad = mn | delta;                                              // negative minimum or negative delta
    if (NN && (delta - mn) != (int)(short)P2) bad = -1;                // delta wrapped around int16: the shortcut does not apply
    B = ((unsigned)delta & 0xffffu) | ((unsigned)bad & 0x80000000u);
}


// ------------------------------------------------------------------------------------------------
// Winner-take-all on packed S (A.5 first half): same result as wta_column, fewer instructions.
//   argmin: one CREDUX.MIN on (S << 16 | d); uniqueness as an order-free predicate; S[best-1], S[best+1] fetched from
//   their owner lanes by shuffle; the small truncating division done in float with an exact +-1 correction.
template <int NWD, bool HASPAD = true>
__device__ __forceinline__ void wta_column2(const unsigned (&s)[NWD], int lane, const SgbmDev& p, int16_t* rawPix, int2* auxPix)
{
    (void)rawPix;
    const int d0 = lane * NWD * 2;
    int key = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < NWD; j++)
        if (!HASPAD || d0 + 2 * j < p.D) {                            // a packed word is either all real disparities or all padding
            key = min(key, (int)((s[j] << 16) | (unsigned)(d0 + 2 * j)));
            key = min(key, (int)((s[j] & 0xffff0000u) | (unsigned)(d0 + 2 * j + 1)));
        }
    key = __reduce_min_sync(0xffffffffu, key);
    const int minS = key >> 16, best = key & 0xffff;
    const int lim = minS * 100, mul = 100 - p.uniq;
    const int rel = best - d0 + 1;                                    // position i is within one of best iff (unsigned)(rel - i) < 3
    bool bad = false;
#pragma unroll
    for (int j = 0; j < NWD; j++) {
        if (!HASPAD || d0 + 2 * j < p.D) {
            const int v0 = lo16(s[j]), v1 = hi16(s[j]);
            bad |= (v0 * mul < lim) && ((unsigned)(rel - 2 * j) >= 3u);
            bad |= (v1 * mul < lim) && ((unsigned)(rel - 2 * j - 1) >= 3u);
        }
    }
    const bool reject = __any_sync(0xffffffffu, bad) || (minS == CVR_MAX_COST);
    int2 aux = make_int2(0x7fffffff, 0);
    if (!reject) {                                                    // warp-uniform
        unsigned ac = 0;
        if (0 < best && best < p.D - 1) {                             // fetch S[best-1], S[best+1] from their owner lanes
            const int ia = best - 1, ic = best + 1;
            unsigned wa = s[0], wc = s[0];
#pragma unroll
            for (int j = 1; j < NWD; j++) { if (((ia >> 1) % NWD) == j) wa = s[j]; if (((ic >> 1) % NWD) == j) wc = s[j]; }
            wa = __shfl_sync(0xffffffffu, wa, ia / (2 * NWD));
            wc = __shfl_sync(0xffffffffu, wc, ic / (2 * NWD));
            const unsigned a = (ia & 1) ? (wa >> 16) : (wa & 0xffffu), c = (ic & 1) ? (wc >> 16) : (wc & 0xffffu);
            ac = (a << 16) | c;
        }
        aux = make_int2((int)(((unsigned)minS << 16) | (unsigned)best), (int)ac);
    }
    if (lane == 0) *auxPix = aux;
}

// keeps a value in a register across the loop (stops the compiler from re-deriving it from constant memory every iteration)
#define CVR_PIN(x) asm volatile("" : "+r"(x))
#define CVR_PIN64(x) asm volatile("" : "+l"(x))

// ------------------------------------------------------------------------------------------------
// Horizontal path, packed.  Same contract as k_path_h.
// Frame selection of a launch (bits 8.. of the `storeSel` argument): 0 all frames; 1 only frames whose regime flags are clear
// when the kernel starts (fast regime: all path costs >= 0 so far, the saturating sums of S do not depend on their order);
// 2 only frames with a flag set.  MODE_HH uses it to run its second pass in the cheaper order (see launch_paths).
__device__ __forceinline__ bool path_frame_skipped(int sel, int flag) { return (sel == 1 && flag != 0) || (sel == 2 && flag == 0); }

template <int DPL, bool REVERSE, bool WTA, bool HASPAD, bool NN>
__global__ void __launch_bounds__(128) k_path_h2(const int16_t* __restrict__ C, int16_t* __restrict__ S, int16_t* __restrict__ raw,
                                                 int2* __restrict__ aux, SgbmDev p, int B, int accumulate, int storeSel, int* flags)
{
    constexpr int NWD = DPL / 2;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (warp >= B * p.H) return;
    const int b = warp / p.H, y = warp % p.H;
    const int storeS = storeSel & 255;
    if (path_frame_skipped(storeSel >> 8, flags[b])) return;
    const int W1 = p.W1, D = p.D;
    int P2 = p.P2;
    const long long colW = (long long)p.Dp / 2;                             // 32-bit words per column
    const size_t rowBase = (((size_t)b * p.H + y) * W1) * colW + (size_t)lane * NWD;
    const int dx = REVERSE ? -1 : 1;
    const long long stepW = dx * colW;
    const unsigned* Cp = (const unsigned*)C + rowBase + (REVERSE ? (size_t)(W1 - 1) * colW : 0);
    unsigned* Sp = (unsigned*)S + rowBase + (REVERSE ? (size_t)(W1 - 1) * colW : 0);
    int16_t* rawp = raw + ((size_t)b * p.H + y) * p.W + p.minX1 + (REVERSE ? W1 - 1 : 0);
    int2* auxp = aux + ((size_t)b * p.H + y) * W1 + (REVERSE ? W1 - 1 : 0);
    int* flagp = flags + b;
    bool exact = *flagp != 0;
    PathK k; k.P1p = pk2(p.P1s); k.Qp = pk2(32767 - p.P1s); k.Z = (unsigned)p.D >> 20; CVR_PIN(k.Z);   // 0 (D < 2^15), but not a compile-time constant
    CVR_PIN(k.P1p); CVR_PIN(k.Qp); CVR_PIN(P2); CVR_PIN(lane);
    unsigned Lp[NWD];
#pragma unroll
    for (int j = 0; j < NWD; j++) Lp[j] = ((lane * NWD + j) * 2 < D) ? 0u : 0x7fff7fffu;
    unsigned iA, iB;
    path_info<NN>(0, P2, iA, iB);                                          // the border column: minimum 0
    __shared__ __align__(16) unsigned ringsm[4][CVR_PF_DEPTH][2][32 * NWD];
    unsigned (*ring)[2][32 * NWD] = ringsm[threadIdx.x >> 5];
    // Prefetch roles: a column of C (and of S) is Dp*2 bytes = CH 16-byte chunks.  Dp <= 128: lanes [0, CH) fetch C, lanes
    // [CH, 2*CH) fetch S with ONE cp.async per lane and column; wider columns: lanes [0, CH) fetch both arrays.
    constexpr int CH = 8 * NWD;
    constexpr bool ONE = (2 * CH <= 32);
    const bool roleS = ONE && lane >= CH;
    const int chunk = roleS ? lane - CH : lane;
    const bool pfOn = ONE ? (lane < (accumulate ? 2 * CH : CH)) : (lane < CH);
    const long long laneOff = (long long)chunk * 4 - (long long)lane * NWD;                     // from this lane's words to its chunk
    const unsigned* pf0 = (roleS ? (const unsigned*)Sp : Cp) + laneOff;                         // next column to fetch (array 0 of this lane)
    const unsigned* pf1 = (const unsigned*)Sp + laneOff;                                        // !ONE: the S stream
    const unsigned ringBase = (unsigned)__cvta_generic_to_shared(&ring[0][roleS ? 1 : 0][chunk * 4]);
    constexpr unsigned slotBytes = 2u * 32u * NWD * 4u, arrBytes = 32u * NWD * 4u;
#pragma unroll
    for (int i = 0; i < CVR_PF_DEPTH; i++) {                               // prologue: the first CVR_PF_DEPTH columns
        if (i < W1 && pfOn) {
            cpa_16(ringBase + i * slotBytes, pf0);
            if (!ONE && accumulate) cpa_16(ringBase + i * slotBytes + arrBytes, pf1);
        }
        cpa_commit();
        pf0 += stepW; pf1 += stepW;
    }
    for (int step = 0; step < W1; step++) {
        unsigned c[NWD], s[NWD], L[NWD];
        const int slot = step % CVR_PF_DEPTH;
        cpa_wait<CVR_PF_DEPTH - 1>();
        __syncwarp();                                                  // the chunks were copied by other lanes
        ldw<NWD>(&ring[slot][0][lane * NWD], c);
        if (accumulate) ldw<NWD>(&ring[slot][1][lane * NWD], s);
        else {
#pragma unroll
            for (int j = 0; j < NWD; j++) s[j] = 0u;
        }
        const bool fastCol = !exact && (int)iB >= 0;
        int m;
        if (fastCol) m = col_fast<NWD, HASPAD, NN>(Lp, iA, c, L, lane, D, k);
        else if (!exact) m = col_general<NWD, false, HASPAD>(Lp, (int)(short)(iB & 0xffffu), c, L, lane, D, k);
        else m = col_general<NWD, true, HASPAD>(Lp, (int)(short)(iB & 0xffffu), c, L, lane, D, k);
        const int mn = __reduce_min_sync(0xffffffffu, m);
        path_info<NN>(mn, P2, iA, iB);
        if (accumulate) {
            if (fastCol && mn >= 0) acc_fast<NWD>(s, L);
            else acc_exact<NWD>(s, L);
        }
        if (mn < 0) {                                                  // negative path cost: later unsigned sums would be void -> exact from here on
            exact = true;
            if (lane == 0) atomicOr(flagp, 2);
        }
#pragma unroll
        for (int j = 0; j < NWD; j++) { Lp[j] = L[j]; if (!accumulate) s[j] = L[j]; }
        if (!WTA || storeS) stw<NWD>(Sp, s);
        if (WTA) wta_column2<NWD, HASPAD>(s, lane, p, rawp, auxp);
        if (step + CVR_PF_DEPTH < W1 && pfOn) {                          // refill the slot just consumed (every lane read it before the reduce)
            cpa_16(ringBase + slot * slotBytes, pf0);
            if (!ONE && accumulate) cpa_16(ringBase + slot * slotBytes + arrBytes, pf1);
        }
        cpa_commit();
        pf0 += stepW; pf1 += stepW;
        Cp += stepW; Sp += stepW; rawp += dx; auxp += dx;
    }
}

// ------------------------------------------------------------------------------------------------
// Vertical/diagonal paths, packed, with the row state in shared memory.
//
// A cluster of CL CTAs owns one frame.  All GW = CL*NW warps of the cluster split the W1 columns into contiguous
// chunks; a warp keeps the previous-row costs of its chunk (3 directions x Dp int16 per column, plus the 3 minima)
// in its CTA's shared memory and updates them IN PLACE, walking its chunk left to right:
//   down-right path : needs the old value of column x'-1  -> carried in registers from the previous iteration
//   down path       : needs the old value of column x'    -> read before it is overwritten
//   down-left path  : needs the old value of column x'+1  -> not yet overwritten when walking left to right
// Only the two chunk-edge columns are needed by other warps; the owner pushes the new edge costs into the
// neighbour's double-buffered inbox (a plain shared-memory store, or a DSMEM store when the neighbour warp lives in
// the next CTA of the cluster).  One cluster barrier per image row orders the pushes against next row's reads.
// HBM traffic is the algorithmic minimum: read C, read S, write S.
struct VState {                // shared-memory carve-up (in 32-bit words) computed on the host
    int colsMax;               // columns per CTA (upper bound)
    int colWords;              // words per column: 3 * (Dp/2 + 4) (k_path_v2 keeps its three minima after the 3 * Dp/2 cost words)
    int inboxWords;            // words per inbox slot: Dp/2 + 4
    int sel;                   // frame selection of this launch (k_path_v4 only; see path_frame_skipped)
};

template <int DPL, bool UP, bool WTA, bool HASPAD>
__global__ void __launch_bounds__((DPL <= 4 ? 1024 : 512), 1) k_path_v2(const int16_t* __restrict__ C, int16_t* __restrict__ S, int16_t* __restrict__ raw,
                                                     int2* __restrict__ aux, SgbmDev p, int CL, int storeS, int* flags, VState vs)
{
    constexpr int NWD = DPL / 2;
    const int NT = blockDim.x, NW = NT / 32;
    extern __shared__ unsigned smv[];
    cg::cluster_group cluster = cg::this_cluster();
    const int b = blockIdx.y, rank = blockIdx.x;
    const int W1 = p.W1, D = p.D, H = p.H;
    int P2 = p.P2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int GW = CL * NW, gw = rank * NW + warp;
    const int ctaFirst = (int)(((long long)W1 * (rank * NW)) / GW);
    const int a = (int)(((long long)W1 * gw) / GW), e = (int)(((long long)W1 * (gw + 1)) / GW);
    const int dwords = p.Dp / 2;                                      // words per direction per column
    const int colWords = vs.colWords;
    unsigned* state = smv;                                            // [colsMax][colWords]
    unsigned* inbox = smv + (size_t)vs.colsMax * colWords;            // [2 parity][NW][2 sides][inboxWords]
    // ---- zero state: L = 0 for d < D, MAX_COST in the padding; minima 0 (also the never-written frame borders) ----
    for (int i = threadIdx.x; i < vs.colsMax * colWords; i += NT) {
        const int w = i % colWords;
        state[i] = (w < 3 * dwords) ? (((w % dwords) * 2 < D) ? 0u : 0x7fff7fffu) : 0u;
    }
    for (int i = threadIdx.x; i < 2 * NW * 2 * vs.inboxWords; i += NT) {
        const int w = i % vs.inboxWords;
        inbox[i] = (w < dwords) ? ((w * 2 < D) ? 0u : 0x7fff7fffu) : 0u;
    }
    bool exact = flags[b] != 0;
    PathK k; k.P1p = pk2(p.P1s); k.Qp = pk2(32767 - p.P1s); k.Z = (unsigned)p.D >> 20; CVR_PIN(k.Z);   // 0 (D < 2^15), but not a compile-time constant
    CVR_PIN(k.P1p); CVR_PIN(k.Qp); CVR_PIN(P2);
    const size_t colW = (size_t)dwords;
    // where this warp's edge costs go: the right neighbour's LEFT inbox (side 0) gets my last column's down-right
    // costs, the left neighbour's RIGHT inbox (side 1) gets my first column's down-left costs
    unsigned* pushRight = nullptr; unsigned* pushLeft = nullptr;
    if (gw + 1 < GW) {
        const int r2 = (gw + 1) / NW, w2 = (gw + 1) % NW;
        unsigned* loc = inbox + ((size_t)w2 * 2 + 0) * vs.inboxWords;
        pushRight = (r2 == rank) ? loc : cluster.map_shared_rank(loc, r2);
    }
    if (gw > 0) {
        const int r2 = (gw - 1) / NW, w2 = (gw - 1) % NW;
        unsigned* loc = inbox + ((size_t)w2 * 2 + 1) * vs.inboxWords;
        pushLeft = (r2 == rank) ? loc : cluster.map_shared_rank(loc, r2);
    }
    const size_t parityStride = (size_t)NW * 2 * vs.inboxWords;
    const unsigned* myInL = inbox + ((size_t)warp * 2 + 0) * vs.inboxWords;
    const unsigned* myInR = inbox + ((size_t)warp * 2 + 1) * vs.inboxWords;
    unsigned* const col0 = state + (size_t)(a - ctaFirst) * colWords + lane * NWD;    // this lane's words of the chunk's first column
    if (CL > 1) cluster.sync(); else __syncthreads();

    for (int t = 0; t < H; t++) {
        const int y = UP ? H - 1 - t : t;
        const int rd = (t + 1) & 1, wr = t & 1;                       // inbox parity read / written in this row
        const size_t rowBase = (((size_t)b * H + y) * W1 + a) * colW + (size_t)lane * NWD;
        const unsigned* Cp = (const unsigned*)C + rowBase;
        unsigned* Sp = (unsigned*)S + rowBase;
        unsigned carry[NWD];
        int carryMin = 0;
        unsigned cN[NWD], sN[NWD];
        if (a < e) {
            ldw<NWD>(Cp, cN);
            ldw<NWD>(Sp, sN);
            ldw<NWD>(myInL + rd * parityStride + lane * NWD, carry);
            carryMin = (int)myInL[rd * parityStride + dwords];
        }
        unsigned* col = col0;
        for (int xp = a; xp < e; xp++, col += colWords, Cp += colW, Sp += colW) {
            const unsigned* colm = col - lane * NWD + 3 * dwords;      // the three minima of this column
            unsigned c[NWD], s[NWD], s0[NWD];
#pragma unroll
            for (int j = 0; j < NWD; j++) { c[j] = cN[j]; s[j] = s0[j] = sN[j]; }
            const bool lastCol = (xp + 1 == e);
            if (!lastCol) { ldw<NWD>(Cp + colW, cN); ldw<NWD>(Sp + colW, sN); }
            // ---- sources (all read before anything of this column is overwritten) ----
            unsigned Lp1[NWD], Lp2[NWD], Lp3[NWD], L1[NWD], L2[NWD], L3[NWD];
#pragma unroll
            for (int j = 0; j < NWD; j++) Lp1[j] = carry[j];
            const int min1 = carryMin;
            ldw<NWD>(col, carry);                                      // old down-right costs of this column: next column's source
            carryMin = (int)colm[0];
            ldw<NWD>(col + dwords, Lp2);
            const int min2 = (int)colm[1];
            int min3;
            if (!lastCol) { ldw<NWD>(col + colWords + 2 * dwords, Lp3); min3 = (int)colm[colWords + 2]; }
            else { ldw<NWD>(myInR + rd * parityStride + lane * NWD, Lp3); min3 = (int)myInR[rd * parityStride + dwords]; }
            int mn1, mn2, mn3;
            if (!exact) {
                mn1 = column_update<NWD, false, HASPAD>(Lp1, min1, c, L1, lane, D, P2, k);
                mn2 = column_update<NWD, false, HASPAD>(Lp2, min2, c, L2, lane, D, P2, k);
                mn3 = column_update<NWD, false, HASPAD>(Lp3, min3, c, L3, lane, D, P2, k);
                acc_fast<NWD>(s, L1); acc_fast<NWD>(s, L2); acc_fast<NWD>(s, L3);
                if (min(min(mn1, mn2), mn3) < 0) {                     // first negative path cost of this frame: redo the sums exactly, stay exact
                    exact = true;
                    if (lane == 0) atomicOr(&flags[b], 2);
#pragma unroll
                    for (int j = 0; j < NWD; j++) s[j] = s0[j];
                    acc_exact<NWD>(s, L1); acc_exact<NWD>(s, L2); acc_exact<NWD>(s, L3);
                }
            } else {
                mn1 = column_update<NWD, true, HASPAD>(Lp1, min1, c, L1, lane, D, P2, k);
                mn2 = column_update<NWD, true, HASPAD>(Lp2, min2, c, L2, lane, D, P2, k);
                mn3 = column_update<NWD, true, HASPAD>(Lp3, min3, c, L3, lane, D, P2, k);
                if (min(min(mn1, mn2), mn3) < 0 && lane == 0) atomicOr(&flags[b], 2);
                acc_exact<NWD>(s, L1); acc_exact<NWD>(s, L2); acc_exact<NWD>(s, L3);
            }
            stw<NWD>(col, L1); stw<NWD>(col + dwords, L2); stw<NWD>(col + 2 * dwords, L3);
            if (lane < 3) (col - lane * NWD)[3 * dwords + lane] = (unsigned)(lane == 0 ? mn1 : (lane == 1 ? mn2 : mn3));
            if (lastCol && pushRight) {
                stw<NWD>(pushRight + wr * parityStride + lane * NWD, L1);
                if (lane == 0) pushRight[wr * parityStride + dwords] = (unsigned)mn1;
            }
            if (xp == a && pushLeft) {
                stw<NWD>(pushLeft + wr * parityStride + lane * NWD, L3);
                if (lane == 0) pushLeft[wr * parityStride + dwords] = (unsigned)mn3;
            }
            if (!WTA || storeS) stw<NWD>(Sp, s);
            if (WTA) wta_column2<NWD>(s, lane, p, raw + ((size_t)b * H + y) * p.W + xp + p.minX1, aux + ((size_t)b * H + y) * W1 + xp);
        }
        if (CL > 1) cluster.sync(); else __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------
// k_path_v3: as k_path_v2, but without any CTA- or cluster-wide barrier inside the row loop.
// A warp only depends on its two neighbour warps (their chunk-edge costs of the previous row).  The edge hand-over
// uses the distributed-shared-memory producer/consumer mechanism of sm_90+/sm_100: the producer writes the 256-byte
// edge with `st.async ... mbarrier::complete_tx::bytes` into the consumer's inbox (same CTA or another CTA of the
// cluster), which counts the bytes on an mbarrier that lives next to that inbox; the consumer posts the expected
// byte count (`mbarrier.arrive.expect_tx`) and waits on the barrier's phase.  No fence, no MEMBAR, and warps drift
// freely (at most one row apart from their neighbours), so global-load latency of one warp overlaps the arithmetic
// of the others.  Inboxes and barriers are double-buffered by row parity; the k-th use of a slot waits on phase
// parity (k & 1).
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned mapa_u32(unsigned addr, unsigned rank)
{
    unsigned r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void st_async_b32(unsigned raddr, unsigned v, unsigned rbar)
{
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(raddr), "r"(v), "r"(rbar) : "memory");
}
template <int NWD>
__device__ __forceinline__ void st_async_words(unsigned raddr, const unsigned (&v)[NWD], unsigned rbar)
{
    if (NWD == 2) asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1, %2}, [%3];" ::"r"(raddr), "r"(v[0]), "r"(v[1]), "r"(rbar) : "memory");
    else if (NWD == 4) asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(raddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(rbar) : "memory");
    else {
#pragma unroll
        for (int i = 0; i < NWD; i++) st_async_b32(raddr + 4 * i, v[i], rbar);
    }
}

__device__ __forceinline__ void st_async_v2(unsigned raddr, unsigned v0, unsigned v1, unsigned rbar)
{
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1, %2}, [%3];" ::"r"(raddr), "r"(v0), "r"(v1), "r"(rbar) : "memory");
}

// Shared-memory layout: one "block" per (column, direction) = Dp/2 words of packed costs, then the two path_info words
// (A, B) of that pixel's minimum and two pad words (keeps every block 16-byte aligned).  An inbox slot is one such block,
// so "the old down-left block of the column to my right" is read through one pointer whether it lives in my own state or
// in the inbox my right neighbour filled.  The column loop is unrolled by two with alternating register sets for the
// prefetched C/S words and for the carried down-right block, which keeps the loop free of register-to-register moves.
struct V3Ctx {                  // per-warp invariants of k_path_v3's column step (everything lives in registers after inlining)
    int lane, D, P2, ncols, b, y, a, W1, storeS;
    bool hasL, hasR, moreRows;
    unsigned pushL_dst, pushL_bar, pushR_dst, pushR_bar;               // remote inbox slot / barrier of this row's parity
    const unsigned* inR;                                               // my right inbox (previous row's parity)
    const unsigned* rowC; unsigned* rowS; long long nextRow;           // this lane's words of the chunk's first column, row pitch in words
    size_t frameRow0;
    int16_t* raw; int2* aux; int* flags;
    PathK k;
};

// stores of one finished column: the three new blocks, their path_info, the edge pushes, S (and the winner when WTA)
template <int NWD, int FT, int LT, bool WTA, bool HASPAD, bool NN, int BLK, int DW>
__device__ __forceinline__ void v3_finish(const V3Ctx& x, const SgbmDev& p, const int ci, unsigned* const cb, unsigned* const Sp, const bool firstCol,
                                          const bool lastCol, const unsigned (&M1)[NWD], const unsigned (&M2)[NWD], const unsigned (&M3)[NWD],
                                          const unsigned (&s)[NWD], const int mn1, const int mn2, const int mn3)
{
    const int lane = x.lane;
    unsigned* const col = cb + lane * NWD;
    stw<NWD>(col, M1); stw<NWD>(col + BLK, M2); stw<NWD>(col + 2 * BLK, M3);
    // lanes 0..2 derive what the next pixel on each path needs from this pixel's minimum
    unsigned iA, iB;
    path_info<NN>(lane == 0 ? mn1 : (lane == 1 ? mn2 : mn3), x.P2, iA, iB);
    if (lane < 3) *(uint2*)(cb + lane * BLK + DW) = make_uint2(iA, iB);
    // the last row's edges have no consumer (nobody posts expect_tx for them): not pushed, so that no st.async can still be
    // in flight towards a CTA that has already left the kernel
    if (FT != 0 && firstCol && x.hasL && x.moreRows) {                // my first column's down-left block -> left neighbour's right inbox
        st_async_words<NWD>(x.pushL_dst + lane * NWD * 4, M3, x.pushL_bar);
        if (lane == 2) st_async_v2(x.pushL_dst + DW * 4, iA, iB, x.pushL_bar);
    }
    if (LT != 0 && lastCol && x.hasR && x.moreRows) {                 // my last column's down-right block -> right neighbour's left inbox
        st_async_words<NWD>(x.pushR_dst + lane * NWD * 4, M1, x.pushR_bar);
        if (lane == 0) st_async_v2(x.pushR_dst + DW * 4, iA, iB, x.pushR_bar);
    }
    if (!WTA || x.storeS) stw<NWD>(Sp, s);
    if (WTA) wta_column2<NWD, HASPAD>(s, lane, p, x.raw + (x.frameRow0 + x.y) * p.W + x.a + ci + p.minX1, x.aux + (x.frameRow0 + x.y) * x.W1 + x.a + ci);
}

// One column of one row.  FT / LT say whether this call can be the chunk's first / last column: 0 = never, 1 = always,
// 2 = decide at run time.  c, s: this column's C and S words (s is updated); cNx, sNx receive the prefetched words of the
// next column (or of the next row's first column); kin/kiIn: the carried old down-right block of the column to the left;
// kout/kiOut receive this column's old down-right block for the next call.
template <int NWD, int FT, int LT, bool WTA, bool HASPAD, bool NN, int BLK, int DW>
__device__ __forceinline__ void v3_column(const V3Ctx& x, const SgbmDev& p, bool& exact, const int ci, unsigned* const cb, const unsigned* const Cp,
                                          unsigned* const Sp, const unsigned (&c)[NWD], unsigned (&s)[NWD], unsigned (&cNx)[NWD], unsigned (&sNx)[NWD],
                                          const unsigned (&kin)[NWD], const uint2 kiIn, unsigned (&kout)[NWD], uint2& kiOut)
{
    constexpr int colWords = 3 * BLK;
    const int lane = x.lane;
    const bool firstCol = FT == 2 ? (ci == 0) : (FT == 1);
    const bool lastCol = LT == 2 ? (ci + 1 == x.ncols) : (LT == 1);
    // ---- prefetch: the next column of this row, or the first column of the next row ----
    if (LT == 0) { ldw<NWD>(Cp + DW, cNx); ldw<NWD>(Sp + DW, sNx); }
    else {
        const unsigned* nC = lastCol ? x.rowC + x.nextRow : Cp + DW;
        const unsigned* nS = lastCol ? x.rowS + x.nextRow : Sp + DW;
        if (!lastCol || x.moreRows) { ldw<NWD>(nC, cNx); ldw<NWD>(nS, sNx); }
    }
    // ---- sources: everything is read before this column's blocks are overwritten ----
    unsigned* const col = cb + lane * NWD;
    unsigned Lp2[NWD], Lp3[NWD], L1[NWD], L2[NWD], L3[NWD];
    if (LT == 0 || (LT == 2 && !lastCol)) { ldw<NWD>(col, kout); kiOut = *(const uint2*)(cb + DW); }   // old down-right block: the next column's source
    ldw<NWD>(col + BLK, Lp2);
    const uint2 i2 = *(const uint2*)(cb + BLK + DW);
    const unsigned* src3 = (LT == 0 || !lastCol) ? cb + colWords + 2 * BLK : x.inR;
    ldw<NWD>(src3 + lane * NWD, Lp3);
    const uint2 i3 = *(const uint2*)(src3 + DW);
    if (!exact && (int)(kiIn.y | i2.y | i3.y) >= 0) {                  // the common case: one straight-line block for all three directions
        const unsigned q1 = col_fast_pk<NWD, HASPAD, NN>(kin, kiIn.x, c, L1, lane, x.D, x.k);
        const unsigned q2 = col_fast_pk<NWD, HASPAD, NN>(Lp2, i2.x, c, L2, lane, x.D, x.k);
        const unsigned q3 = col_fast_pk<NWD, HASPAD, NN>(Lp3, i3.x, c, L3, lane, x.D, x.k);
        // lane minima of two directions side by side; a 32-bit signed minimum over the warp is decided by the high half
        const unsigned z12 = __vmins2(__byte_perm(q1, q2, 0x5410), __byte_perm(q1, q2, 0x7632));   // (min1 | min2 << 16)
        const unsigned z3 = __vmins2(q3, __byte_perm(q3, q3, 0x1032));                              // min3 in both halves
        const int r2 = __reduce_min_sync(0xffffffffu, (int)z12);
        const int r1 = __reduce_min_sync(0xffffffffu, (int)(z12 << 16));
        const int r3 = __reduce_min_sync(0xffffffffu, (int)z3);
        const int mn1 = r1 >> 16, mn2 = r2 >> 16, mn3 = r3 >> 16;
        if ((r1 | r2 | r3) < 0) {                                      // a negative path cost: unsigned sums are void, stay exact from here on
            exact = true;
            if (lane == 0) atomicOr(&x.flags[x.b], 2);
            acc_exact<NWD>(s, L1); acc_exact<NWD>(s, L2); acc_exact<NWD>(s, L3);
        } else {
            acc_fast<NWD>(s, L1); acc_fast<NWD>(s, L2); acc_fast<NWD>(s, L3);
        }
        v3_finish<NWD, FT, LT, WTA, HASPAD, NN, BLK, DW>(x, p, ci, cb, Sp, firstCol, lastCol, L1, L2, L3, s, mn1, mn2, mn3);
    } else {
        const int d1 = (int)(short)(kiIn.y & 0xffffu), d2 = (int)(short)(i2.y & 0xffffu), d3 = (int)(short)(i3.y & 0xffffu);
        int m1, m2, m3;
        if (!exact) {
            m1 = col_general<NWD, false, HASPAD>(kin, d1, c, L1, lane, x.D, x.k);
            m2 = col_general<NWD, false, HASPAD>(Lp2, d2, c, L2, lane, x.D, x.k);
            m3 = col_general<NWD, false, HASPAD>(Lp3, d3, c, L3, lane, x.D, x.k);
        } else {
            m1 = col_general<NWD, true, HASPAD>(kin, d1, c, L1, lane, x.D, x.k);
            m2 = col_general<NWD, true, HASPAD>(Lp2, d2, c, L2, lane, x.D, x.k);
            m3 = col_general<NWD, true, HASPAD>(Lp3, d3, c, L3, lane, x.D, x.k);
        }
        const int mn1 = __reduce_min_sync(0xffffffffu, m1);
        const int mn2 = __reduce_min_sync(0xffffffffu, m2);
        const int mn3 = __reduce_min_sync(0xffffffffu, m3);
        if (min(min(mn1, mn2), mn3) < 0 && !exact) {
            exact = true;
            if (lane == 0) atomicOr(&x.flags[x.b], 2);
        }
        acc_exact<NWD>(s, L1); acc_exact<NWD>(s, L2); acc_exact<NWD>(s, L3);   // valid for any signs
        v3_finish<NWD, FT, LT, WTA, HASPAD, NN, BLK, DW>(x, p, ci, cb, Sp, firstCol, lastCol, L1, L2, L3, s, mn1, mn2, mn3);
    }
}

template <int DPL, bool UP, bool WTA, bool HASPAD, bool NN, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) k_path_v3(const int16_t* __restrict__ C, int16_t* __restrict__ S, int16_t* __restrict__ raw,
                                                     int2* __restrict__ aux, SgbmDev p, int CL, int storeS, int* flags, VState vs)
{
    constexpr int NWD = DPL / 2;
    const int NT = blockDim.x, NW = NT / 32;
    extern __shared__ uint4 smv3[];
    cg::cluster_group cluster = cg::this_cluster();
    const int b = blockIdx.y, rank = blockIdx.x;
    const int W1 = p.W1, D = p.D, H = p.H;
    int P2 = p.P2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int GW = CL * NW, gw = rank * NW + warp;
    const int ctaFirst = (int)(((long long)W1 * (rank * NW)) / GW);
    const int a = (int)(((long long)W1 * gw) / GW), e = (int)(((long long)W1 * (gw + 1)) / GW);
    constexpr int dwords = DPL * 16;                                  // Dp / 2 words per direction per column (Dp = 32 * DPL)
    constexpr int blk = dwords + 4;                                   // costs, A, B (path_info), 2 pad words
    constexpr int colWords = 3 * blk;
    unsigned* const state = (unsigned*)smv3;                          // [colsMax][3 directions][blk]
    unsigned* const inbox = state + (size_t)vs.colsMax * colWords;    // [2 parity][NW][2 sides][blk]
    const size_t parityStride = (size_t)NW * 2 * blk;
    unsigned long long* bars = (unsigned long long*)(inbox + 2 * parityStride);
    {
        unsigned A0, B0;
        path_info<NN>(0, P2, A0, B0);                                 // the zero state: all costs 0 (padding: MAX_COST), minimum 0
        const int nwords = (vs.colsMax * 3 + 2 * NW * 2) * blk;       // state and inboxes are one run of blocks
        for (int i = threadIdx.x; i < nwords; i += NT) {
            const int w = i % blk;
            state[i] = (w < dwords) ? ((w * 2 < D) ? 0u : 0x7fff7fffu) : (w == dwords ? A0 : (w == dwords + 1 ? B0 : 0u));
        }
    }
    if (threadIdx.x < 2 * NW * 2) mbar_init(smem_u32(bars + threadIdx.x), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    bool exact = flags[b] != 0;
    PathK k; k.P1p = pk2(p.P1s); k.Qp = pk2(32767 - p.P1s); k.Z = (unsigned)p.D >> 20; CVR_PIN(k.Z);   // 0 (D < 2^15), but not a compile-time constant
    CVR_PIN(k.P1p); CVR_PIN(k.Qp); CVR_PIN(P2);
    constexpr size_t colW = (size_t)dwords;
    const unsigned edgeBytes = 32u * NWD * 4u + 8u;                   // 32 lanes x NWD words + A + B
    const unsigned* myInL = inbox + ((size_t)warp * 2 + 0) * blk;
    const unsigned* myInR = inbox + ((size_t)warp * 2 + 1) * blk;
    const unsigned myBarL = smem_u32(bars + (size_t)warp * 2 + 0), myBarR = smem_u32(bars + (size_t)warp * 2 + 1);
    const unsigned barParityStride = (unsigned)(NW * 2 * 8), inboxParityBytes = (unsigned)(parityStride * 4);
    const bool hasL = gw > 0, hasR = gw + 1 < GW;
    unsigned pushR_data = 0, pushR_bar = 0, pushL_data = 0, pushL_bar = 0;
    if (hasR) {
        const int r2 = (gw + 1) / NW, w2 = (gw + 1) % NW;
        pushR_data = mapa_u32(smem_u32(inbox + ((size_t)w2 * 2 + 0) * blk), r2);
        pushR_bar = mapa_u32(smem_u32(bars + (size_t)w2 * 2 + 0), r2);
    }
    if (hasL) {
        const int r2 = (gw - 1) / NW, w2 = (gw - 1) % NW;
        pushL_data = mapa_u32(smem_u32(inbox + ((size_t)w2 * 2 + 1) * blk), r2);
        pushL_bar = mapa_u32(smem_u32(bars + (size_t)w2 * 2 + 1), r2);
    }
    unsigned* const colBase0 = state + (size_t)(a - ctaFirst) * colWords;
    cluster.sync();                                                   // state, inboxes and barriers of every CTA are initialised

    const size_t frameRow0 = (size_t)b * H;
    const long long nextRow = (UP ? -1LL : 1LL) * (long long)W1 * (long long)colW;
    const size_t rb0 = ((frameRow0 + (UP ? H - 1 : 0)) * W1 + a) * colW + (size_t)lane * NWD;
    const unsigned* rowC = (const unsigned*)C + rb0;                  // this lane's words of column a in the current row
    unsigned* rowS = (unsigned*)S + rb0;
    const int ncols = e - a;
    V3Ctx x;
    x.lane = lane; x.D = D; x.P2 = P2; x.ncols = ncols; x.b = b; x.a = a; x.W1 = W1; x.storeS = storeS; x.hasL = hasL; x.hasR = hasR;
    x.nextRow = nextRow; x.frameRow0 = frameRow0; x.raw = raw; x.aux = aux; x.flags = flags; x.k = k;
    unsigned cA[NWD], sA[NWD], cB[NWD], sB[NWD];                       // C and S words of the current / the next column (roles alternate)
    ldw<NWD>(rowC, cA);
    ldw<NWD>(rowS, sA);
    for (int t = 0; t < H; t++, rowC += nextRow, rowS += nextRow) {
        const int y = UP ? H - 1 - t : t;
        const int rd = (t + 1) & 1, wr = t & 1;
        if (t > 0) {                                                   // previous row's edges of both neighbours
            const unsigned waitPhase = (unsigned)(((t - 1) >> 1) & 1);
            if (lane == 0) {
                if (hasL) mbar_arrive_expect_tx(myBarL + rd * barParityStride, edgeBytes);
                if (hasR) mbar_arrive_expect_tx(myBarR + rd * barParityStride, edgeBytes);
            }
            __syncwarp();                                              // also: the info words lanes 0..2 wrote in the previous row
            if (hasL) mbar_wait(myBarL + rd * barParityStride, waitPhase);
            if (hasR) mbar_wait(myBarR + rd * barParityStride, waitPhase);
        }
        const unsigned* inL = myInL + rd * parityStride;               // filled by the neighbours in the previous row; this row's
        const unsigned* inR = myInR + rd * parityStride;               // pushes go to the other parity
        x.y = y; x.moreRows = t + 1 < H; x.inR = inR; x.rowC = rowC; x.rowS = rowS;
        x.pushL_dst = pushL_data + wr * inboxParityBytes; x.pushL_bar = pushL_bar + wr * barParityStride;
        x.pushR_dst = pushR_data + wr * inboxParityBytes; x.pushR_bar = pushR_bar + wr * barParityStride;
        unsigned kA[NWD], kB[NWD];                                     // the carried down-right block (costs) ...
        uint2 kiA, kiB;                                                // ... and its (A, B)
        ldw<NWD>(inL + lane * NWD, kA);
        kiA = *(const uint2*)(inL + dwords);

        unsigned* cb = colBase0;                                       // base of the current column's blocks
        const unsigned* Cp = rowC;
        unsigned* Sp = rowS;
        int ci = 0;
        for (; ci + 2 <= ncols; ci += 2, cb += 2 * colWords, Cp += 2 * colW, Sp += 2 * colW) {
            v3_column<NWD, 2, 0, WTA, HASPAD, NN, blk, dwords>(x, p, exact, ci, cb, Cp, Sp, cA, sA, cB, sB, kA, kiA, kB, kiB);
            v3_column<NWD, 0, 2, WTA, HASPAD, NN, blk, dwords>(x, p, exact, ci + 1, cb + colWords, Cp + colW, Sp + colW, cB, sB, cA, sA, kB, kiB, kA, kiA);
        }
        if (ci < ncols) {                                              // odd chunk: one more column, then hand the prefetched words back
            v3_column<NWD, 2, 1, WTA, HASPAD, NN, blk, dwords>(x, p, exact, ci, cb, Cp, Sp, cA, sA, cB, sB, kA, kiA, kB, kiB);
#pragma unroll
            for (int j = 0; j < NWD; j++) { cA[j] = cB[j]; sA[j] = sB[j]; }
        }
    }
    cluster.sync();                                                   // nobody exits while a neighbour may still write into its inbox
}

// ------------------------------------------------------------------------------------------------
// k_path_h3: the horizontal path with SEVERAL ROWS PER WARP.
// k_path_h2 gives a warp one row, a lane DPL disparities: at D = 128 the packed arithmetic of a column is a dozen
// instructions while the warp-uniform rest (loop, addresses, prefetch, the minimum over the column, path_info, the scalar
// half of winner-take-all) is ten times that -- the kernel is bound by instruction issue, not by HBM.  Here a group of LPR
// lanes owns a row (Dp / LPR disparities per lane, 32 / LPR rows per warp), so that per-column overhead is shared by all
// rows of the warp; minima are taken over the lane group with xor-shuffles.
// Ownership is PIECE-INTERLEAVED: a column is cut into NP pieces of LPR * PW words; lane `sub` of the group owns words
// [sub * PW, sub * PW + PW) of every piece.  One cp.async / one store instruction of the group then covers LPR * PW * 4
// contiguous bytes (whole 32-byte sectors; a lane-contiguous split requested every sector twice and ran into the L2 read
// bandwidth), and every lane prefetches exactly the words it consumes itself, so cp.async.wait_group is the only
// synchronisation of the ring (layout [array][piece][lane][PW]: conflict-free 16-byte accesses).  Arithmetic and regimes
// (fast / no-neighbour / general / exact) are those of k_path_h2; a column takes the fast form only when every row of
// the warp may.
template <int NWD, int LPR, int PW>
__device__ __forceinline__ int h3_word(int sub, int j) { return (j / PW) * (LPR * PW) + sub * PW + (j % PW); }   // column word index of this lane's word j

// neighbour words of the lane's words: lm1[j] = (L[2w-1], L[2w]), lp1[j] = (L[2w+1], L[2w+2]) as packed pairs
template <int NWD, int LPR, int PW>
__device__ __forceinline__ void h3_neighbours(const unsigned (&Lp)[NWD], int sub, unsigned (&lm1)[NWD], unsigned (&lp1)[NWD])
{
    constexpr int NP = NWD / PW;
#pragma unroll
    for (int q = 0; q < NP; q++) {
        // word before this lane's first word of piece q: the last word of piece q of lane sub-1, or of piece q-1 of the last lane
        unsigned up = __shfl_up_sync(0xffffffffu, Lp[q * PW + PW - 1], 1, LPR);
        const unsigned wrapUp = q > 0 ? __shfl_sync(0xffffffffu, Lp[(q > 0 ? q - 1 : 0) * PW + PW - 1], LPR - 1, LPR) : 0x7fff7fffu;
        if (sub == 0) up = wrapUp;
        unsigned dn = __shfl_down_sync(0xffffffffu, Lp[q * PW], 1, LPR);
        const unsigned wrapDn = q + 1 < NP ? __shfl_sync(0xffffffffu, Lp[(q + 1 < NP ? q + 1 : q) * PW], 0, LPR) : 0x7fff7fffu;
        if (sub == LPR - 1) dn = wrapDn;
#pragma unroll
        for (int e = 0; e < PW; e++) {
            const int j = q * PW + e;
            lm1[j] = __byte_perm(e == 0 ? up : Lp[j - 1], Lp[j], 0x5432);
            lp1[j] = __byte_perm(Lp[j], e == PW - 1 ? dn : Lp[j + 1], 0x5432);
        }
    }
}
template <int NWD, int LPR, int PW, bool HASPAD, bool NN>
__device__ __forceinline__ unsigned colx_fast_pk(const unsigned (&Lp)[NWD], unsigned ndeltap, const unsigned (&c)[NWD], unsigned (&L)[NWD], int sub, int D, const PathK& k)
{
    unsigned mnp = 0x7fff7fffu;
    if (NN) {
#pragma unroll
        for (int j = 0; j < NWD; j++) {
            unsigned v = __vadd2(__viaddmin_s16x2(Lp[j], ndeltap, k.Z), c[j]);
            if (HASPAD && h3_word<NWD, LPR, PW>(sub, j) * 2 >= D) v = 0x7fff7fffu;
            L[j] = v;
            mnp = __vmins2(mnp, v);
        }
    } else {
        unsigned lm1[NWD], lp1[NWD];
        h3_neighbours<NWD, LPR, PW>(Lp, sub, lm1, lp1);
#pragma unroll
        for (int j = 0; j < NWD; j++) {
            const unsigned n = __vimin3_s16x2(lm1[j], lp1[j], k.Qp);
            const unsigned t = __viaddmin_s16x2(n, k.P1p, Lp[j]);
            unsigned v = __vadd2(__viaddmin_s16x2(t, ndeltap, k.Z), c[j]);
            if (HASPAD && h3_word<NWD, LPR, PW>(sub, j) * 2 >= D) v = 0x7fff7fffu;
            L[j] = v;
            mnp = __vmins2(mnp, v);
        }
    }
    return mnp;
}
template <int NWD, int LPR, int PW, bool EXACT, bool HASPAD>
__device__ __forceinline__ unsigned colx_general_pk(const unsigned (&Lp)[NWD], int delta, const unsigned (&c)[NWD], unsigned (&L)[NWD], int sub, int D, const PathK& k)
{
    unsigned lm1[NWD], lp1[NWD];
    h3_neighbours<NWD, LPR, PW>(Lp, sub, lm1, lp1);
    const unsigned deltap = pk2(delta), lop = pk2(max(delta - 32768, -32768)), ndeltap = pk2(-delta);
    unsigned mnp = 0x7fff7fffu;
#pragma unroll
    for (int j = 0; j < NWD; j++) {
        unsigned v = path_cost2<EXACT>(lm1[j], Lp[j], lp1[j], deltap, lop, ndeltap, k, c[j]);
        if (HASPAD && h3_word<NWD, LPR, PW>(sub, j) * 2 >= D) v = 0x7fff7fffu;
        L[j] = v;
        mnp = __vmins2(mnp, v);
    }
    return mnp;
}
// minimum over the LPR lanes of a row group (every lane of the group receives it)
template <int LPR>
__device__ __forceinline__ int group_min(int v)
{
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

#define CVR_H3_DEPTH 3          // columns in flight per warp
#define CVR_H3_WARPS 4          // warps per CTA
template <int NWD, int LPR, bool REVERSE, bool WTA, bool HASPAD, bool NN>
__global__ void __launch_bounds__(CVR_H3_WARPS * 32) k_path_h3(const int16_t* __restrict__ C, int16_t* __restrict__ S, int2* __restrict__ aux, SgbmDev p, int B,
                                                               int accumulate, int storeSel, int* flags)
{
    const int storeS = storeSel & 255;
    constexpr int R = 32 / LPR;                                      // rows per warp
    constexpr int PW = (NWD % 4 == 0) ? 4 : 2;                       // words per piece (16- or 8-byte accesses)
    constexpr int NP = NWD / PW;                                     // pieces per column
    constexpr int PIECEW = LPR * PW;                                 // words per piece of a column
    constexpr int SLOTW = 2 * NWD * 32;                              // words per ring slot: [array][piece][lane][PW]
    __shared__ __align__(16) unsigned ringsm[CVR_H3_WARPS][CVR_H3_DEPTH][SLOTW];
    const int wib = threadIdx.x >> 5;
    int lane = threadIdx.x & 31;
    const int wpf = (p.H + R - 1) / R;                               // warps per frame: the rows of a warp belong to one frame
    const int gw = blockIdx.x * CVR_H3_WARPS + wib;
    if (gw >= B * wpf) return;
    const int b = gw / wpf;
    if (path_frame_skipped(storeSel >> 8, flags[b])) return;
    const int sub = lane % LPR, rsel = lane / LPR;
    const int y0 = (gw - b * wpf) * R + rsel;
    const bool rowOn = y0 < p.H;                                     // a group past the last row shadows it (same loads, no stores)
    const int y = rowOn ? y0 : p.H - 1;
    const int W1 = p.W1, D = p.D;
    int P2 = p.P2;
    const long long colW = (long long)p.Dp / 2;                      // 32-bit words per column
    const size_t rowBase = (((size_t)b * p.H + y) * W1) * colW + (size_t)sub * PW;
    const int dx = REVERSE ? -1 : 1;
    const long long stepW = dx * colW;
    const unsigned* Cp = (const unsigned*)C + rowBase + (REVERSE ? (size_t)(W1 - 1) * colW : 0);   // this lane's words of piece 0 of the current column
    unsigned* Sp = (unsigned*)S + rowBase + (REVERSE ? (size_t)(W1 - 1) * colW : 0);
    int2* auxp = aux + ((size_t)b * p.H + y) * W1 + (REVERSE ? W1 - 1 : 0);
    int* flagp = flags + b;
    bool exact = *flagp != 0;
    PathK k; k.P1p = pk2(p.P1s); k.Qp = pk2(32767 - p.P1s); k.Z = (unsigned)p.D >> 20; CVR_PIN(k.Z);
    CVR_PIN(k.P1p); CVR_PIN(k.Qp); CVR_PIN(P2); CVR_PIN(lane);
    unsigned Lp[NWD];
#pragma unroll
    for (int j = 0; j < NWD; j++) Lp[j] = (h3_word<NWD, LPR, PW>(sub, j) * 2 < D) ? 0u : 0x7fff7fffu;
    unsigned iA, iB;
    path_info<NN>(0, P2, iA, iB);                                    // the border column: minimum 0
    unsigned* const ring = &ringsm[wib][0][0];
    const unsigned ringBase = (unsigned)__cvta_generic_to_shared(ring) + (unsigned)lane * PW * 4u;
    const unsigned* pfC = Cp;
    const unsigned* pfS = Sp;
    auto prefetch = [&](int slot) {
#pragma unroll
        for (int q = 0; q < NP; q++) {
            const unsigned dst = ringBase + (unsigned)(slot * SLOTW + q * 32 * PW) * 4u;
            if (PW == 4) {
                cpa_16(dst, pfC + q * PIECEW);
                if (accumulate) cpa_16(dst + NP * 32 * PW * 4u, pfS + q * PIECEW);
            } else {
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(pfC + q * PIECEW) : "memory");
                if (accumulate) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst + NP * 32 * PW * 4u), "l"(pfS + q * PIECEW) : "memory");
            }
        }
    };
#pragma unroll
    for (int i = 0; i < CVR_H3_DEPTH; i++) {
        if (i < W1) prefetch(i);
        cpa_commit();
        pfC += stepW; pfS += stepW;
    }
    const int mul = 100 - p.uniq;
    int slot = 0;
    for (int step = 0; step < W1; step++) {
        unsigned c[NWD], s[NWD], L[NWD];
        cpa_wait<CVR_H3_DEPTH - 1>();                                 // this lane's own pieces of the oldest slot have landed
        const unsigned* rs = ring + slot * SLOTW + lane * PW;
#pragma unroll
        for (int q = 0; q < NP; q++) {
            if (PW == 4) { const uint4 a = *(const uint4*)(rs + q * 32 * PW); c[4 * q] = a.x; c[4 * q + 1] = a.y; c[4 * q + 2] = a.z; c[4 * q + 3] = a.w; }
            else { const uint2 a = *(const uint2*)(rs + q * 32 * PW); c[2 * q] = a.x; c[2 * q + 1] = a.y; }
        }
        if (accumulate) {
#pragma unroll
            for (int q = 0; q < NP; q++) {
                if (PW == 4) { const uint4 a = *(const uint4*)(rs + (NP + q) * 32 * PW); s[4 * q] = a.x; s[4 * q + 1] = a.y; s[4 * q + 2] = a.z; s[4 * q + 3] = a.w; }
                else { const uint2 a = *(const uint2*)(rs + (NP + q) * 32 * PW); s[2 * q] = a.x; s[2 * q + 1] = a.y; }
            }
        } else {
#pragma unroll
            for (int j = 0; j < NWD; j++) s[j] = 0u;
        }
        const bool fastCol = !exact && __all_sync(0xffffffffu, (int)iB >= 0);     // every row of the warp is in the fast regime
        unsigned mnp;
        if (fastCol) mnp = colx_fast_pk<NWD, LPR, PW, HASPAD, NN>(Lp, iA, c, L, sub, D, k);
        else if (!exact) mnp = colx_general_pk<NWD, LPR, PW, false, HASPAD>(Lp, (int)(short)(iB & 0xffffu), c, L, sub, D, k);
        else mnp = colx_general_pk<NWD, LPR, PW, true, HASPAD>(Lp, (int)(short)(iB & 0xffffu), c, L, sub, D, k);
        const int mn = group_min<LPR>(min((int)(short)(mnp & 0xffffu), ((int)mnp) >> 16));
        path_info<NN>(mn, P2, iA, iB);
        const bool anyNeg = __any_sync(0xffffffffu, mn < 0);
        if (accumulate) {
            if (fastCol && !anyNeg) acc_fast<NWD>(s, L);
            else acc_exact<NWD>(s, L);
        }
        if (anyNeg) {                                                // negative path cost: later unsigned sums would be void -> exact from here on
            exact = true;
            if (mn < 0 && sub == 0) atomicOr(flagp, 2);
        }
#pragma unroll
        for (int j = 0; j < NWD; j++) { Lp[j] = L[j]; if (!accumulate) s[j] = L[j]; }
        if ((!WTA || storeS) && rowOn) {
#pragma unroll
            for (int q = 0; q < NP; q++) {
                if (PW == 4) *(uint4*)(Sp + q * PIECEW) = make_uint4(s[4 * q], s[4 * q + 1], s[4 * q + 2], s[4 * q + 3]);
                else *(uint2*)(Sp + q * PIECEW) = make_uint2(s[2 * q], s[2 * q + 1]);
            }
        }
        if (WTA) {
            // A.5 first half for the row of this lane group (as wta_column2)
            int key = 0x7fffffff;
#pragma unroll
            for (int j = 0; j < NWD; j++) {
                const int d = h3_word<NWD, LPR, PW>(sub, j) * 2;
                if (!HASPAD || d < D) {
                    key = min(key, (int)((s[j] << 16) | (unsigned)d));
                    key = min(key, (int)((s[j] & 0xffff0000u) | (unsigned)(d + 1)));
                }
            }
            key = group_min<LPR>(key);
            const int minS = key >> 16, best = key & 0xffff;
            const int lim = minS * 100;
            bool bad = false;
#pragma unroll
            for (int j = 0; j < NWD; j++) {
                const int d = h3_word<NWD, LPR, PW>(sub, j) * 2;
                if (!HASPAD || d < D) {
                    const int v0 = lo16(s[j]), v1 = hi16(s[j]);
                    bad |= (v0 * mul < lim) && ((unsigned)(best + 1 - d) >= 3u);     // |d - best| > 1
                    bad |= (v1 * mul < lim) && ((unsigned)(best - d) >= 3u);         // |d + 1 - best| > 1
                }
            }
            // the updated S of the column goes back into the slot's S pieces (same conflict-free layout), where the leader lane of
            // each row picks up S[best - 1] and S[best + 1]
            unsigned* ws = ring + slot * SLOTW + (NP * 32 + lane) * PW;
#pragma unroll
            for (int q = 0; q < NP; q++) {
                if (PW == 4) *(uint4*)(ws + q * 32 * PW) = make_uint4(s[4 * q], s[4 * q + 1], s[4 * q + 2], s[4 * q + 3]);
                else *(uint2*)(ws + q * 32 * PW) = make_uint2(s[2 * q], s[2 * q + 1]);
            }
            __syncwarp();
            const unsigned badMask = __ballot_sync(0xffffffffu, bad);
            const unsigned grp = (LPR == 32 ? 0xffffffffu : ((1u << LPR) - 1u)) << (rsel * LPR);
            const bool reject = (badMask & grp) != 0u || minS == CVR_MAX_COST;
            if (sub == 0 && rowOn) {
                int2 a2 = make_int2(0x7fffffff, 0);
                if (!reject) {
                    unsigned ac = 0;
                    if (0 < best && best < D - 1) {
                        auto fetch = [&](int d) -> unsigned {         // S of disparity d of this row, from the slot
                            const int w = d >> 1, q = w / PIECEW, r = w - q * PIECEW, so = r / PW, e = r - so * PW;
                            const unsigned v = ring[slot * SLOTW + ((NP + q) * 32 + rsel * LPR + so) * PW + e];
                            return (d & 1) ? (v >> 16) : (v & 0xffffu);
                        };
                        ac = (fetch(best - 1) << 16) | fetch(best + 1);
                    }
                    a2 = make_int2((int)(((unsigned)minS << 16) | (unsigned)best), (int)ac);
                }
                *auxp = a2;
            }
            __syncwarp();                                            // the slot is refilled below
        }
        if (step + CVR_H3_DEPTH < W1) prefetch(slot);                 // refill the slot just consumed (only this lane ever reads these pieces)
        cpa_commit();
        pfC += stepW; pfS += stepW;
        Cp += stepW; Sp += stepW; auxp += dx;
        slot = (slot + 1 == CVR_H3_DEPTH) ? 0 : slot + 1;
    }
}

// ------------------------------------------------------------------------------------------------
// k_path_v4: the vertical / diagonal paths with SEVERAL COLUMNS PER WARP.
// Same cluster-per-frame organisation, shared-memory row state, inboxes and st.async / mbarrier edge hand-over as
// k_path_v3.  The columns of a row do not depend on each other (all three paths read the PREVIOUS row), so a warp can work
// on G = 32 / LPR adjacent columns of its chunk at once: a group of LPR lanes owns a column (Dp / LPR disparities per
// lane, piece-interleaved like k_path_h3 so that every global access of a group covers whole sectors).  The warp-uniform
// part of a column step (loop, addresses, the minima over the column, path_info, edge logic) -- two thirds of
// k_path_v3's instructions at D = 128 -- is then shared by G columns.  The state is still updated in place: a step first
// reads everything it needs (old blocks of columns x-1 .. x+G), then writes the G new columns; only the down-right
// block of the step's last column is needed again by the next step (its first column's source) and is carried in the
// registers of group 0.  512 threads per CTA (the register budget of 40 packed words of state per lane).
template <int NWD, int LPR, int PW, bool HASPAD>
__device__ __forceinline__ void group_wta(const unsigned (&s)[NWD], int sub, int grp, int D, int uniq, bool on, int2* auxPix)
{
    constexpr int PIECEW = LPR * PW;
    int key = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < NWD; j++) {
        const int d = h3_word<NWD, LPR, PW>(sub, j) * 2;
        if (!HASPAD || d < D) {
            key = min(key, (int)((s[j] << 16) | (unsigned)d));
            key = min(key, (int)((s[j] & 0xffff0000u) | (unsigned)(d + 1)));
        }
    }
    key = group_min<LPR>(key);
    const int minS = key >> 16, best = key & 0xffff;
    const int lim = minS * 100, mul = 100 - uniq;
    bool bad = false;
#pragma unroll
    for (int j = 0; j < NWD; j++) {
        const int d = h3_word<NWD, LPR, PW>(sub, j) * 2;
        if (!HASPAD || d < D) {
            const int v0 = lo16(s[j]), v1 = hi16(s[j]);
            bad |= (v0 * mul < lim) && ((unsigned)(best + 1 - d) >= 3u);
            bad |= (v1 * mul < lim) && ((unsigned)(best - d) >= 3u);
        }
    }
    const unsigned badMask = __ballot_sync(0xffffffffu, bad);
    const unsigned gm = (LPR == 32 ? 0xffffffffu : ((1u << LPR) - 1u)) << (grp * LPR);
    const bool reject = (badMask & gm) != 0u || minS == CVR_MAX_COST;
    // S[best - 1], S[best + 1] from their owner lanes (every lane takes part in the shuffles)
    const bool inner = 0 < best && best < D - 1;
    unsigned va = 0, vc = 0;
    {
        const int da = inner ? best - 1 : 0, dc = inner ? best + 1 : 0;
        const int wa = da >> 1, wc = dc >> 1;
        const int qa = wa / PIECEW, ra = wa - qa * PIECEW, soa = ra / PW, ja = qa * PW + (ra - soa * PW);
        const int qc = wc / PIECEW, rc = wc - qc * PIECEW, soc = rc / PW, jc = qc * PW + (rc - soc * PW);
        unsigned ma = s[0], mc = s[0];
#pragma unroll
        for (int j = 1; j < NWD; j++) { if (ja == j) ma = s[j]; if (jc == j) mc = s[j]; }
        ma = __shfl_sync(0xffffffffu, ma, grp * LPR + soa);
        mc = __shfl_sync(0xffffffffu, mc, grp * LPR + soc);
        va = (da & 1) ? (ma >> 16) : (ma & 0xffffu);
        vc = (dc & 1) ? (mc >> 16) : (mc & 0xffffu);
    }
    if (sub == 0 && on) {
        int2 a2 = make_int2(0x7fffffff, 0);
        if (!reject) a2 = make_int2((int)(((unsigned)minS << 16) | (unsigned)best), inner ? (int)((va << 16) | vc) : 0);
        *auxPix = a2;
    }
}
// packed 16-bit minimum over the lanes of a group, both halves at once
template <int LPR>
__device__ __forceinline__ unsigned group_min2(unsigned v)
{
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) v = __vmins2(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

template <int NWD, int LPR, bool UP, bool WTA, bool HASPAD, bool NN, int MAXT = 512>
__global__ void __launch_bounds__(MAXT, 1) k_path_v4(const int16_t* __restrict__ C, int16_t* __restrict__ S, int16_t* __restrict__ raw,
                                                    int2* __restrict__ aux, SgbmDev p, int CL, int storeS, int* flags, VState vs)
{
    (void)raw;
    // (every CTA of the cluster decides alike: a frame's flag only changes after the first cluster.sync of a kernel working on it;
    // the selection travels in the parameter struct, not in a register: the kernel sits at its 128-register limit)
    bool exact = flags[blockIdx.y] != 0;
    if (vs.sel != 0 && (vs.sel == 1) == exact) asm volatile("exit;");   // (a plain `return` here costs the kernel three spilled registers)
    constexpr int G = 32 / LPR;                                       // columns per step
    constexpr int PW = (NWD % 4 == 0) ? 4 : 2, NP = NWD / PW, PIECEW = LPR * PW;
    constexpr int dwords = NWD * LPR;                                 // Dp / 2 words per direction per column
    constexpr int blk = dwords + 4;                                   // costs, A, B (path_info), 2 pad words
    constexpr int colWords = 3 * blk;
    const int NT = blockDim.x, NW = NT / 32;
    extern __shared__ uint4 smv4[];
    cg::cluster_group cluster = cg::this_cluster();
    const int b = blockIdx.y, rank = blockIdx.x;
    const int W1 = p.W1, D = p.D, H = p.H;
    int P2 = p.P2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sub = lane % LPR, g = lane / LPR;
    const int GW = CL * NW, gw = rank * NW + warp;
    const int ctaFirst = (int)(((long long)W1 * (rank * NW)) / GW);
    const int a = (int)(((long long)W1 * gw) / GW), e = (int)(((long long)W1 * (gw + 1)) / GW);
    unsigned* const state = (unsigned*)smv4;                          // [colsMax][3 directions][blk]
    unsigned* const inbox = state + (size_t)vs.colsMax * colWords;    // [2 parity][NW][2 sides][blk]
    const size_t parityStride = (size_t)NW * 2 * blk;
    unsigned long long* bars = (unsigned long long*)(inbox + 2 * parityStride);
    {
        unsigned A0, B0;
        path_info<NN>(0, P2, A0, B0);                                 // the zero state: all costs 0 (padding: MAX_COST), minimum 0
        const int nwords = (vs.colsMax * 3 + 2 * NW * 2) * blk;       // state and inboxes are one run of blocks
        for (int i = threadIdx.x; i < nwords; i += NT) {
            const int w = i % blk;
            state[i] = (w < dwords) ? ((w * 2 < D) ? 0u : 0x7fff7fffu) : (w == dwords ? A0 : (w == dwords + 1 ? B0 : 0u));
        }
    }
    if (threadIdx.x < 2 * NW * 2) mbar_init(smem_u32(bars + threadIdx.x), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    PathK k; k.P1p = pk2(p.P1s); k.Qp = pk2(32767 - p.P1s); k.Z = (unsigned)p.D >> 20; CVR_PIN(k.Z);
    CVR_PIN(k.P1p); CVR_PIN(k.Qp); CVR_PIN(P2);
    constexpr size_t colW = (size_t)dwords;
    const unsigned edgeBytes = (unsigned)dwords * 4u + 8u;             // a block's costs + A + B
    const unsigned* myInL = inbox + ((size_t)warp * 2 + 0) * blk;
    const unsigned* myInR = inbox + ((size_t)warp * 2 + 1) * blk;
    const unsigned myBarL = smem_u32(bars + (size_t)warp * 2 + 0), myBarR = smem_u32(bars + (size_t)warp * 2 + 1);
    const unsigned barParityStride = (unsigned)(NW * 2 * 8), inboxParityBytes = (unsigned)(parityStride * 4);
    const bool hasL = gw > 0, hasR = gw + 1 < GW;
    unsigned pushR_data = 0, pushR_bar = 0, pushL_data = 0, pushL_bar = 0;
    if (hasR) {
        const int r2 = (gw + 1) / NW, w2 = (gw + 1) % NW;
        pushR_data = mapa_u32(smem_u32(inbox + ((size_t)w2 * 2 + 0) * blk), r2);
        pushR_bar = mapa_u32(smem_u32(bars + (size_t)w2 * 2 + 0), r2);
    }
    if (hasL) {
        const int r2 = (gw - 1) / NW, w2 = (gw - 1) % NW;
        pushL_data = mapa_u32(smem_u32(inbox + ((size_t)w2 * 2 + 1) * blk), r2);
        pushL_bar = mapa_u32(smem_u32(bars + (size_t)w2 * 2 + 1), r2);
    }
    unsigned* const colBase0 = state + (size_t)(a - ctaFirst) * colWords;
    cluster.sync();                                                   // state, inboxes and barriers of every CTA are initialised

    const int ncols = e - a;
    const size_t frameRow0 = (size_t)b * H;
    const long long nextRow = (UP ? -1LL : 1LL) * (long long)W1 * (long long)colW;
    // this lane's words of piece 0 of column a in the current row
    const size_t rb0 = ((frameRow0 + (UP ? H - 1 : 0)) * W1 + a) * colW + (size_t)sub * PW;
    const unsigned* rowC = (const unsigned*)C + rb0;
    unsigned* rowS = (unsigned*)S + rb0;
    const int lw = sub * PW;                                          // this lane's word offset inside a piece
    auto ldblk = [&](const unsigned* blkp, unsigned (&v)[NWD]) {      // this lane's words of a block in shared memory
#pragma unroll
        for (int q = 0; q < NP; q++) {
            if (PW == 4) { const uint4 t4 = *(const uint4*)(blkp + q * PIECEW + lw); v[4 * q] = t4.x; v[4 * q + 1] = t4.y; v[4 * q + 2] = t4.z; v[4 * q + 3] = t4.w; }
            else { const uint2 t2 = *(const uint2*)(blkp + q * PIECEW + lw); v[2 * q] = t2.x; v[2 * q + 1] = t2.y; }
        }
    };
    auto stblk = [&](unsigned* blkp, const unsigned (&v)[NWD]) {
#pragma unroll
        for (int q = 0; q < NP; q++) {
            if (PW == 4) *(uint4*)(blkp + q * PIECEW + lw) = make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
            else *(uint2*)(blkp + q * PIECEW + lw) = make_uint2(v[2 * q], v[2 * q + 1]);
        }
    };
    auto ldg = [&](const unsigned* gp, unsigned (&v)[NWD]) {          // this lane's words of a column in global memory
#pragma unroll
        for (int q = 0; q < NP; q++) {
            if (PW == 4) { const uint4 t4 = *(const uint4*)(gp + q * PIECEW); v[4 * q] = t4.x; v[4 * q + 1] = t4.y; v[4 * q + 2] = t4.z; v[4 * q + 3] = t4.w; }
            else { const uint2 t2 = *(const uint2*)(gp + q * PIECEW); v[2 * q] = t2.x; v[2 * q + 1] = t2.y; }
        }
    };
    // the first step's C and S words (column a + g, clamped into the chunk for groups past its end)
    unsigned cN[NWD], sN[NWD];
    {
        const int c0 = min(g, ncols - 1);
        ldg(rowC + (size_t)c0 * colW, cN);
        ldg(rowS + (size_t)c0 * colW, sN);
    }
    for (int t = 0; t < H; t++, rowC += nextRow, rowS += nextRow) {
        const int y = UP ? H - 1 - t : t;
        const int rd = (t + 1) & 1, wr = t & 1;
        const bool moreRows = t + 1 < H;
        if (t > 0) {                                                   // previous row's edges of both neighbours
            const unsigned waitPhase = (unsigned)(((t - 1) >> 1) & 1);
            if (lane == 0) {
                if (hasL) mbar_arrive_expect_tx(myBarL + rd * barParityStride, edgeBytes);
                if (hasR) mbar_arrive_expect_tx(myBarR + rd * barParityStride, edgeBytes);
            }
            __syncwarp();                                              // also: the blocks and info words written in the previous row
            if (hasL) mbar_wait(myBarL + rd * barParityStride, waitPhase);
            if (hasR) mbar_wait(myBarR + rd * barParityStride, waitPhase);
        }
        const unsigned* inL = myInL + rd * parityStride;               // filled by the neighbours in the previous row; this row's
        const unsigned* inR = myInR + rd * parityStride;               // pushes go to the other parity
        const unsigned pL_dst = pushL_data + wr * inboxParityBytes, pL_bar = pushL_bar + wr * barParityStride;
        const unsigned pR_dst = pushR_data + wr * inboxParityBytes, pR_bar = pushR_bar + wr * barParityStride;
        unsigned carry[NWD];                                           // group 0: old down-right block of the column left of this step
        uint2 carryI;
        ldblk(inL, carry);
        carryI = *(const uint2*)(inL + dwords);
        // this group's column pointers, advanced by G columns per step (the group's column is clamped into the chunk only in a
        // part-filled last step)
        const unsigned* gC = rowC + (size_t)g * colW;
        unsigned* gS = rowS + (size_t)g * colW;
        for (int ci = 0; ci < ncols; ci += G, gC += G * colW, gS += G * colW) {
            const int col = ci + g;
            const bool on = col < ncols;
            const bool lastStep = ci + G >= ncols;
            const int cc = min(col, ncols - 1);                        // groups past the chunk's end shadow its last column (no stores)
            unsigned* const cb = colBase0 + cc * colWords;
            unsigned c[NWD], s[NWD];
#pragma unroll
            for (int j = 0; j < NWD; j++) { c[j] = cN[j]; s[j] = sN[j]; }
            // ---- prefetch: the next step of this row, or the first step of the next row ----
            if (!lastStep) {
                const long long back = (long long)(col + G - min(col + G, ncols - 1)) * (long long)colW;   // 0 unless the next step is part-filled
                ldg(gC + G * colW - back, cN); ldg(gS + G * colW - back, sN);
            } else if (moreRows) {
                const size_t o = (size_t)min(g, ncols - 1) * colW;
                ldg(rowC + nextRow + o, cN); ldg(rowS + nextRow + o, sN);
            }
            // ---- sources: everything is read before any block of this step is overwritten ----
            unsigned Lp1[NWD], Lp2[NWD], Lp3[NWD];
            uint2 i1, i2, i3;
            {                                                          // (no divergent branches here: every group does the same loads, group 0 then takes the carried block)
                const unsigned* src1 = cb - (cc > 0 ? colWords : 0);   // old down-right block of column x-1
                ldblk(src1, Lp1);
                i1 = *(const uint2*)(src1 + dwords);
                const bool g0 = g == 0;
#pragma unroll
                for (int j = 0; j < NWD; j++) Lp1[j] = g0 ? carry[j] : Lp1[j];
                i1.x = g0 ? carryI.x : i1.x; i1.y = g0 ? carryI.y : i1.y;
            }
            ldblk(cb + blk, Lp2);
            i2 = *(const uint2*)(cb + blk + dwords);
            const unsigned* src3 = (cc + 1 < ncols) ? cb + colWords + 2 * blk : inR;
            ldblk(src3, Lp3);
            i3 = *(const uint2*)(src3 + dwords);
            {                                                          // the next step's first column reads the old down-right block of this step's last column
                const unsigned* nx = colBase0 + min(ci + G - 1, ncols - 1) * colWords;
                ldblk(nx, carry);
                carryI = *(const uint2*)(nx + dwords);
            }
            __syncwarp();                                              // all reads of the step precede its writes
            unsigned L1[NWD], L2[NWD], L3[NWD];
            unsigned q1, q2, q3;
            const bool fast = !exact && __all_sync(0xffffffffu, (int)(i1.y | i2.y | i3.y) >= 0);
            if (fast) {
                q1 = colx_fast_pk<NWD, LPR, PW, HASPAD, NN>(Lp1, i1.x, c, L1, sub, D, k);
                q2 = colx_fast_pk<NWD, LPR, PW, HASPAD, NN>(Lp2, i2.x, c, L2, sub, D, k);
                q3 = colx_fast_pk<NWD, LPR, PW, HASPAD, NN>(Lp3, i3.x, c, L3, sub, D, k);
            } else {
                const int d1 = (int)(short)(i1.y & 0xffffu), d2 = (int)(short)(i2.y & 0xffffu), d3 = (int)(short)(i3.y & 0xffffu);
                if (!exact) {
                    q1 = colx_general_pk<NWD, LPR, PW, false, HASPAD>(Lp1, d1, c, L1, sub, D, k);
                    q2 = colx_general_pk<NWD, LPR, PW, false, HASPAD>(Lp2, d2, c, L2, sub, D, k);
                    q3 = colx_general_pk<NWD, LPR, PW, false, HASPAD>(Lp3, d3, c, L3, sub, D, k);
                } else {
                    q1 = colx_general_pk<NWD, LPR, PW, true, HASPAD>(Lp1, d1, c, L1, sub, D, k);
                    q2 = colx_general_pk<NWD, LPR, PW, true, HASPAD>(Lp2, d2, c, L2, sub, D, k);
                    q3 = colx_general_pk<NWD, LPR, PW, true, HASPAD>(Lp3, d3, c, L3, sub, D, k);
                }
            }
            // lane minima of two directions side by side, reduced over the group as packed pairs
            const unsigned z12 = group_min2<LPR>(__vmins2(__byte_perm(q1, q2, 0x5410), __byte_perm(q1, q2, 0x7632)));   // (min1 | min2 << 16)
            const unsigned z3 = group_min2<LPR>(__vmins2(q3, __byte_perm(q3, q3, 0x1032)));                              // min3 in both halves
            const bool anyNeg = __any_sync(0xffffffffu, (int)((z12 | z3) & 0x80008000u) != 0);
            if (fast && !anyNeg) { acc_fast<NWD>(s, L1); acc_fast<NWD>(s, L2); acc_fast<NWD>(s, L3); }
            else { acc_exact<NWD>(s, L1); acc_exact<NWD>(s, L2); acc_exact<NWD>(s, L3); }       // valid for any signs
            if (anyNeg && !exact) {                                    // a negative path cost: unsigned sums are void, stay exact from here on
                exact = true;
                if (lane == 0) atomicOr(&flags[b], 2);
            }
            // ---- stores of the finished columns ----
            // lanes 0..2 of the group derive what the next pixel on "their" path needs from this pixel's minimum
            const int mnSel = sub == 0 ? (int)(short)(z12 & 0xffffu) : (sub == 1 ? (int)z12 >> 16 : (int)z3 >> 16);
            unsigned iA, iB;
            path_info<NN>(mnSel, P2, iA, iB);
            if (on) {
                stblk(cb, L1); stblk(cb + blk, L2); stblk(cb + 2 * blk, L3);
                if (sub < 3) *(uint2*)(cb + sub * blk + dwords) = make_uint2(iA, iB);
                // the last row's edges have no consumer: not pushed (no st.async may be in flight towards a CTA that has left)
                if (moreRows && (col == 0 || lastStep)) {
                    if (col == 0 && hasL) {                            // my first column's down-left block -> left neighbour's right inbox
#pragma unroll
                        for (int q = 0; q < NP; q++) {
                            if (PW == 4) asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(pL_dst + (q * PIECEW + lw) * 4), "r"(L3[4 * q]), "r"(L3[4 * q + 1]), "r"(L3[4 * q + 2]), "r"(L3[4 * q + 3]), "r"(pL_bar) : "memory");
                            else st_async_v2(pL_dst + (q * PIECEW + lw) * 4, L3[2 * q], L3[2 * q + 1], pL_bar);
                        }
                        if (sub == 2) st_async_v2(pL_dst + dwords * 4, iA, iB, pL_bar);
                    }
                    if (col == ncols - 1 && hasR) {                    // my last column's down-right block -> right neighbour's left inbox
#pragma unroll
                        for (int q = 0; q < NP; q++) {
                            if (PW == 4) asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(pR_dst + (q * PIECEW + lw) * 4), "r"(L1[4 * q]), "r"(L1[4 * q + 1]), "r"(L1[4 * q + 2]), "r"(L1[4 * q + 3]), "r"(pR_bar) : "memory");
                            else st_async_v2(pR_dst + (q * PIECEW + lw) * 4, L1[2 * q], L1[2 * q + 1], pR_bar);
                        }
                        if (sub == 0) st_async_v2(pR_dst + dwords * 4, iA, iB, pR_bar);
                    }
                }
                if (!WTA || storeS) {
#pragma unroll
                    for (int q = 0; q < NP; q++) {
                        if (PW == 4) *(uint4*)(gS + q * PIECEW) = make_uint4(s[4 * q], s[4 * q + 1], s[4 * q + 2], s[4 * q + 3]);
                        else *(uint2*)(gS + q * PIECEW) = make_uint2(s[2 * q], s[2 * q + 1]);
                    }
                }
            }
            if (WTA) group_wta<NWD, LPR, PW, HASPAD>(s, sub, g, D, p.uniq, on, aux + (frameRow0 + y) * W1 + a + cc);
        }
    }
    cluster.sync();                                                   // nobody exits while a neighbour may still write into its inbox
}
