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
};

template <bool P1NEG>
__device__ __forceinline__ unsigned path_cost2(unsigned lm1, unsigned l0, unsigned lp1, unsigned deltap, unsigned lop, unsigned ndeltap,
                                               const PathK& k, unsigned c, bool cneg)
{
    unsigned n;
    if (!P1NEG) n = __vimin3_s16x2(lm1, lp1, k.Qp);
    else n = __vmaxs2(__vmins2(lm1, lp1), k.Qp);
    const unsigned t = __viaddmin_s16x2(n, k.P1p, l0);               // min(n + P1s, Lp[d])
    const unsigned m = __vmaxs2(__vmins2(t, deltap), lop);           // min(., delta), kept >= delta - 32768
    const unsigned t2 = __vadd2(m, ndeltap);                         // m - delta in [-32768, 0]
    if (!cneg) return __vadd2(t2, c);
    const unsigned w = __vadd2(~t2, 0x80018001u);                    // -32768 - t2
    return __vadd2(__vmaxs2(c, w), t2);                              // max(t2 + C, -32768)
}

__device__ __forceinline__ unsigned acc_s2(unsigned s, unsigned L, bool sneg)
{
    return sneg ? __vaddss2(s, L) : __viaddmin_u16x2(s, L, 0x7fff7fffu);
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

// One column update for one direction: Lp (previous costs of the source pixel, packed) -> L (this pixel), returns
// the warp-wide minimum of L.  `left`/`right` are the neighbouring lanes' boundary words (fetched by the caller).
template <int NWD, bool P1NEG, bool HASPAD>
__device__ __forceinline__ int column_update(const unsigned (&Lp)[NWD], int minPrev, const unsigned (&c)[NWD], unsigned (&L)[NWD], int lane, int D,
                                             int P2, const PathK& k, bool cneg)
{
    const int delta = (int)(short)(P2 + minPrev);
    const unsigned deltap = pk2(delta), lop = pk2(max(delta - 32768, -32768)), ndeltap = pk2(-delta);
    unsigned left = __shfl_up_sync(0xffffffffu, Lp[NWD - 1], 1);
    unsigned right = __shfl_down_sync(0xffffffffu, Lp[0], 1);
    if (lane == 0) left = 0x7fff7fffu;                 // Lp[-1] = MAX_COST
    if (lane == 31) right = 0x7fff7fffu;               // Lp[D]  = MAX_COST (padding lanes hold MAX_COST too)
    unsigned mnp = 0x7fff7fffu;
#pragma unroll
    for (int j = 0; j < NWD; j++) {
        const unsigned lm1 = __byte_perm(j == 0 ? left : Lp[j - 1], Lp[j], 0x5432);           // (Lp[2j-1], Lp[2j])
        const unsigned lp1 = __byte_perm(Lp[j], j == NWD - 1 ? right : Lp[j + 1], 0x5432);     // (Lp[2j+1], Lp[2j+2])
        unsigned v = path_cost2<P1NEG>(lm1, Lp[j], lp1, deltap, lop, ndeltap, k, c[j], cneg);
        if (HASPAD && (lane * NWD + j) * 2 >= D) v = 0x7fff7fffu;
        L[j] = v;
        mnp = __vmins2(mnp, v);
    }
    const int m = min((int)(short)(mnp & 0xffffu), ((int)mnp) >> 16);
    return __reduce_min_sync(0xffffffffu, m);
}

// ------------------------------------------------------------------------------------------------
// Horizontal path, packed.  Same contract as k_path_h.
template <int DPL, bool REVERSE, bool WTA, bool HASPAD>
__global__ void __launch_bounds__(128) k_path_h2(const int16_t* __restrict__ C, int16_t* __restrict__ S, int16_t* __restrict__ raw,
                                                 int* __restrict__ aux, SgbmDev p, int B, int accumulate, int storeS, int* flags)
{
    constexpr int NWD = DPL / 2;
    constexpr bool P1NEG = false;            // P1s < 0 is routed to the int32 kernels by the host
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= B * p.H) return;
    const int b = warp / p.H, y = warp % p.H;
    const int W1 = p.W1, D = p.D;
    const size_t colW = (size_t)p.Dp / 2;                                   // 32-bit words per column
    const size_t rowBase = (((size_t)b * p.H + y) * W1) * colW + (size_t)lane * NWD;
    const unsigned* Crow = (const unsigned*)C + rowBase;
    unsigned* Srow = (unsigned*)S + rowBase;
    const int fl = flags[b];
    const bool cneg = (fl & 1) != 0;
    bool sneg = (fl & 2) != 0;
    PathK k; k.P1p = pk2(p.P1s); k.Qp = P1NEG ? pk2(-32768 - p.P1s) : pk2(32767 - p.P1s);
    unsigned Lp[NWD];
#pragma unroll
    for (int j = 0; j < NWD; j++) Lp[j] = ((lane * NWD + j) * 2 < D) ? 0u : 0x7fff7fffu;
    int minPrev = 0;
    int xp = REVERSE ? W1 - 1 : 0;
    const int dx = REVERSE ? -1 : 1;
    unsigned cN[NWD], sN[NWD];
#pragma unroll
    for (int j = 0; j < NWD; j++) sN[j] = 0u;
    ldw<NWD>(Crow + (size_t)xp * colW, cN);
    if (accumulate) ldw<NWD>(Srow + (size_t)xp * colW, sN);
    for (int step = 0; step < W1; step++, xp += dx) {
        unsigned c[NWD], s[NWD], L[NWD];
#pragma unroll
        for (int j = 0; j < NWD; j++) { c[j] = cN[j]; s[j] = sN[j]; }
        if (step + 1 < W1) {
            ldw<NWD>(Crow + (size_t)(xp + dx) * colW, cN);
            if (accumulate) ldw<NWD>(Srow + (size_t)(xp + dx) * colW, sN);
        }
        minPrev = column_update<NWD, P1NEG, HASPAD>(Lp, minPrev, c, L, lane, D, p.P2, k, cneg);
        if (minPrev < 0 && !sneg) { sneg = true; if (lane == 0) atomicOr(&flags[b], 2); }
#pragma unroll
        for (int j = 0; j < NWD; j++) { Lp[j] = L[j]; s[j] = accumulate ? acc_s2(s[j], L[j], sneg) : L[j]; }
        if (!WTA || storeS) stw<NWD>(Srow + (size_t)xp * colW, s);
        if (WTA) {
            int Si[DPL];
#pragma unroll
            for (int j = 0; j < NWD; j++) { Si[2 * j] = lo16(s[j]); Si[2 * j + 1] = hi16(s[j]); }
            wta_column<DPL>(Si, lane, p, raw + ((size_t)b * p.H + y) * p.W + xp + p.minX1, aux + ((size_t)b * p.H + y) * W1 + xp);
        }
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
    int colWords;              // words per column: 3 * Dp/2 + 4 (three minima + pad)
    int inboxWords;            // words per inbox slot: Dp/2 + 4
};

template <int DPL, bool UP, bool WTA, bool HASPAD>
__global__ void __launch_bounds__(1024, 1) k_path_v2(const int16_t* __restrict__ C, int16_t* __restrict__ S, int16_t* __restrict__ raw,
                                                   int* __restrict__ aux, SgbmDev p, int CL, int storeS, int* flags, VState vs)
{
    constexpr int NWD = DPL / 2;
    constexpr bool P1NEG = false;            // P1s < 0 is routed to the int32 kernels by the host
    const int NT = blockDim.x, NW = NT / 32;
    extern __shared__ unsigned smv[];
    cg::cluster_group cluster = cg::this_cluster();
    const int b = blockIdx.y, rank = blockIdx.x;
    const int W1 = p.W1, D = p.D, H = p.H;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int GW = CL * NW, gw = rank * NW + warp;
    const int ctaFirst = (int)(((long long)W1 * (rank * NW)) / GW);
    const int a = (int)(((long long)W1 * gw) / GW), e = (int)(((long long)W1 * (gw + 1)) / GW);
    const int dwords = p.Dp / 2;                                      // words per direction per column
    unsigned* state = smv;                                            // [colsMax][colWords]
    unsigned* inbox = smv + (size_t)vs.colsMax * vs.colWords;         // [2 parity][NW][2 sides][inboxWords]
    // ---- zero state: L = 0 for d < D, MAX_COST in the padding; minima 0 (also the never-written frame borders) ----
    for (int i = threadIdx.x; i < vs.colsMax * vs.colWords; i += NT) {
        const int w = i % vs.colWords;
        state[i] = (w < 3 * dwords) ? (((w % dwords) * 2 < D) ? 0u : 0x7fff7fffu) : 0u;
    }
    for (int i = threadIdx.x; i < 2 * NW * 2 * vs.inboxWords; i += NT) {
        const int w = i % vs.inboxWords;
        inbox[i] = (w < dwords) ? ((w * 2 < D) ? 0u : 0x7fff7fffu) : 0u;
    }
    const int fl = flags[b];
    const bool cneg = (fl & 1) != 0;
    bool sneg = (fl & 2) != 0;
    PathK k; k.P1p = pk2(p.P1s); k.Qp = P1NEG ? pk2(-32768 - p.P1s) : pk2(32767 - p.P1s);
    const size_t colW = (size_t)dwords;
    // where this warp's edge costs go: right neighbour's LEFT inbox (side 0) gets my last column's down-right costs,
    // left neighbour's RIGHT inbox (side 1) gets my first column's down-left costs
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
    if (CL > 1) cluster.sync(); else __syncthreads();

    for (int t = 0; t < H; t++) {
        const int y = UP ? H - 1 - t : t;
        const int rd = (t + 1) & 1, wr = t & 1;                       // inbox parity read / written in this row
        const size_t rowBase = (((size_t)b * H + y) * W1) * colW + (size_t)lane * NWD;
        const unsigned* Crow = (const unsigned*)C + rowBase;
        unsigned* Srow = (unsigned*)S + rowBase;
        unsigned carry[NWD];
        int carryMin;
        unsigned cN[NWD], sN[NWD];
        if (a < e) {
            ldw<NWD>(Crow + (size_t)a * colW, cN);
            ldw<NWD>(Srow + (size_t)a * colW, sN);
            ldw<NWD>(myInL + rd * parityStride + lane * NWD, carry);
            carryMin = (int)myInL[rd * parityStride + dwords];
        }
        for (int xp = a; xp < e; xp++) {
            unsigned* col = state + (size_t)(xp - ctaFirst) * vs.colWords;
            unsigned c[NWD], s[NWD];
#pragma unroll
            for (int j = 0; j < NWD; j++) { c[j] = cN[j]; s[j] = sN[j]; }
            if (xp + 1 < e) {
                ldw<NWD>(Crow + (size_t)(xp + 1) * colW, cN);
                ldw<NWD>(Srow + (size_t)(xp + 1) * colW, sN);
            }
            unsigned L[NWD], Lp[NWD];
            int mn;
            // ---- down-right (source: column x'-1, previous row) ----
#pragma unroll
            for (int j = 0; j < NWD; j++) Lp[j] = carry[j];
            int minSrc = carryMin;
            ldw<NWD>(col + lane * NWD, carry);                         // old costs of this column: next column's source
            carryMin = (int)col[3 * dwords + 0];
            mn = column_update<NWD, P1NEG, HASPAD>(Lp, minSrc, c, L, lane, D, p.P2, k, cneg);
            if (mn < 0 && !sneg) { sneg = true; if (lane == 0) atomicOr(&flags[b], 2); }
#pragma unroll
            for (int j = 0; j < NWD; j++) s[j] = acc_s2(s[j], L[j], sneg);
            stw<NWD>(col + lane * NWD, L);
            if (lane == 0) col[3 * dwords + 0] = (unsigned)mn;
            if (xp == e - 1 && pushRight) {
                stw<NWD>(pushRight + wr * parityStride + lane * NWD, L);
                if (lane == 0) pushRight[wr * parityStride + dwords] = (unsigned)mn;
            }
            // ---- down (source: this column, previous row) ----
            ldw<NWD>(col + dwords + lane * NWD, Lp);
            minSrc = (int)col[3 * dwords + 1];
            mn = column_update<NWD, P1NEG, HASPAD>(Lp, minSrc, c, L, lane, D, p.P2, k, cneg);
            if (mn < 0 && !sneg) { sneg = true; if (lane == 0) atomicOr(&flags[b], 2); }
#pragma unroll
            for (int j = 0; j < NWD; j++) s[j] = acc_s2(s[j], L[j], sneg);
            stw<NWD>(col + dwords + lane * NWD, L);
            if (lane == 0) col[3 * dwords + 1] = (unsigned)mn;
            // ---- down-left (source: column x'+1, previous row) ----
            if (xp + 1 < e) {
                ldw<NWD>(col + vs.colWords + 2 * dwords + lane * NWD, Lp);
                minSrc = (int)col[vs.colWords + 3 * dwords + 2];
            } else {
                ldw<NWD>(myInR + rd * parityStride + lane * NWD, Lp);
                minSrc = (int)myInR[rd * parityStride + dwords];
            }
            mn = column_update<NWD, P1NEG, HASPAD>(Lp, minSrc, c, L, lane, D, p.P2, k, cneg);
            if (mn < 0 && !sneg) { sneg = true; if (lane == 0) atomicOr(&flags[b], 2); }
#pragma unroll
            for (int j = 0; j < NWD; j++) s[j] = acc_s2(s[j], L[j], sneg);
            stw<NWD>(col + 2 * dwords + lane * NWD, L);
            if (lane == 0) col[3 * dwords + 2] = (unsigned)mn;
            if (xp == a && pushLeft) {
                stw<NWD>(pushLeft + wr * parityStride + lane * NWD, L);
                if (lane == 0) pushLeft[wr * parityStride + dwords] = (unsigned)mn;
            }
            __syncwarp();
            if (!WTA || storeS) stw<NWD>(Srow + (size_t)xp * colW, s);
            if (WTA) {
                int Si[DPL];
#pragma unroll
                for (int j = 0; j < NWD; j++) { Si[2 * j] = lo16(s[j]); Si[2 * j + 1] = hi16(s[j]); }
                wta_column<DPL>(Si, lane, p, raw + ((size_t)b * H + y) * p.W + xp + p.minX1, aux + ((size_t)b * H + y) * W1 + xp);
            }
        }
        if (CL > 1) cluster.sync(); else __syncthreads();
    }
}
