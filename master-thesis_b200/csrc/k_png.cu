// k_png.cu -- SURVEY 8(f) row n1, the step BEFORE the path: the two cv::imdecode(message.*ImageData, cv::IMREAD_COLOR) calls of
// MessageConverter::ConvertStereoMessage (cv_reconstruct/src/server/MessageConverter.cpp:65-66).  The streamers send every frame
// as a PNG made by cv::imencode(".png", img) (kitti_vision_streamer/src/main.cpp:217-220, stereopi_video_streamer/src/main.cpp:
// 230-231; wire format network_protocol/src/core/StereoStream.cpp:69-126), so at the rate the path now runs the host would have to
// inflate thousands of PNGs per second.  Here a batch of PNG files is decoded on the device into the BGR frames the path reads:
//   host      : walk the chunk list of each file (signature, IHDR, the IDAT segments) -- byte bookkeeping only
//   k_png_idat: the IDAT payloads of every file -> one contiguous zlib stream per image
//   k_png_inflate: zlib / DEFLATE (stored, fixed and dynamic Huffman blocks), ONE THREAD PER IMAGE (a single-warp CTA each): the
//               bit stream is read and the scanlines are written directly in global memory (the output is its own 32 KiB
//               window), only the Huffman tables live in shared memory (first-level tables, canonical-code search for the
//               rare longer codes) -- 32 streams in flight per SM; a stream is a chain of dependent instructions, so the rate
//               of the decoder is the number of streams in flight (measured: inflate_all below)
//   k_png_unfilter: the five PNG scanline filters (None / Sub / Up / Average / Paeth, sequential per image where the filter
//               demands it, the colour chains of a row in parallel) and the channel order of IMREAD_COLOR (RGB(A) / gray -> BGR)
// Supported: 8-bit, non-interlaced, colour types 0 (gray), 2 (RGB) and 6 (RGBA; alpha dropped as IMREAD_COLOR does).  Anything
// else is refused with CVR_ERR_UNSUPPORTED (the caller keeps its host decoder for such files).  Results are bit-exact against
// cv2.imdecode (tests/test_png.py): PNG is lossless, there is exactly one right answer.
#include "cvr_internal.cuh"
#include <string.h>

struct PngImg {                 // per image, built on the host
    long long fileOff;          // offset of the file in the uploaded byte blob
    int segFirst, segCount;     // its IDAT segments in the segment table
    long long zOff, zLen;       // offset / length of its zlib stream in the gathered buffer
    int colorType, bpp;         // 0 / 2 / 6; bytes per pixel 1 / 3 / 4
    int status;                 // device: 0 ok, else an error code of the inflater
};
struct PngSeg { long long src, dst; int len; int img; };

__global__ void k_png_idat(const uint8_t* __restrict__ blob, const PngSeg* __restrict__ segs, int nseg, uint8_t* __restrict__ z)
{
    for (int s = blockIdx.x; s < nseg; s += gridDim.x) {
        const PngSeg g = segs[s];
        for (int i = threadIdx.x; i < g.len; i += blockDim.x) z[g.dst + i] = blob[g.src + i];
    }
}

// ---- inflate ---------------------------------------------------------------------------------------------------------
#define PNG_LBITS 10            // first-level table bits, literal/length codes
#define PNG_DBITS 8             // ... distance codes
#ifdef __CUDA_ARCH__
#define PNG_BREV(x) __brev(x)
#else
static inline unsigned png_brev_host(unsigned v) { unsigned r = 0; for (int i = 0; i < 32; i++) { r = (r << 1) | (v & 1); v >>= 1; } return r; }
#define PNG_BREV(x) png_brev_host(x)
#endif
struct Huff {                   // canonical code: counts per length, symbols sorted by (length, symbol)
    unsigned short count[16];
    unsigned short symbol[288];
};
struct InflateTab {             // per image: shared memory on the device
    unsigned short ltab[1 << PNG_LBITS];   // (length << 9) | symbol; 0 = code longer than the table
    unsigned short dtab[1 << PNG_DBITS];
    Huff lh, dh;
    unsigned short lens[320];
};

__host__ __device__ inline int huff_build(Huff& h, const unsigned short* lens, int n, unsigned short* tab, int tbits)
{
    for (int i = 0; i < 16; i++) h.count[i] = 0;
    for (int i = 0; i < n; i++) h.count[lens[i]]++;
    if (h.count[0] == n) { for (int i = 0; i < (1 << tbits); i++) tab[i] = 0; return 0; }   // no codes at all (legal for the distance alphabet)
    int left = 1;
    for (int l = 1; l < 16; l++) { left <<= 1; left -= h.count[l]; if (left < 0) return -1; }   // over-subscribed
    unsigned short offs[16];
    offs[1] = 0;
    for (int l = 1; l < 15; l++) offs[l + 1] = offs[l] + h.count[l];
    for (int i = 0; i < n; i++) if (lens[i]) h.symbol[offs[lens[i]]++] = (unsigned short)i;
    // first-level table: every tbits-bit pattern whose low bits are a (bit-reversed) code of length <= tbits
    for (int i = 0; i < (1 << tbits); i++) tab[i] = 0;
    int code = 0, idx = 0;
    for (int l = 1; l <= tbits; l++) {
        for (int k = 0; k < h.count[l]; k++, idx++, code++) {
            unsigned rev = PNG_BREV((unsigned)code) >> (32 - l);
            const unsigned short e = (unsigned short)((l << 9) | h.symbol[idx]);
            for (unsigned f = rev; f < (1u << tbits); f += (1u << l)) tab[f] = e;
        }
        code <<= 1;
    }
    return left > 0 ? 1 : 0;    // incomplete codes are tolerated where zlib tolerates them (single distance code)
}

// zlib / DEFLATE of one stream, start to end, by ONE thread: `in` holds inLen bytes, is 4-byte aligned and readable for 8 bytes
// more (the bit buffer is refilled a 32-bit word at a time); the output goes straight to `out` (rawLen bytes), which also serves
// as the 32 KiB window -- a match reads what this thread wrote earlier.  Nothing of a stream lives in shared memory but its
// Huffman tables (4.4 KB), so that 32 streams are in flight per SM: the decoder is a chain of dependent instructions per symbol
// and its throughput is the number of streams in flight.
// Returns 0 (the Adler-32 is checked by k_png_unfilter: 8), or: 1 bad zlib header, 2 bad block type, 3 bad code lengths, 4 bad symbol / distance, 5 output size differs from
// rawLen, 6 input overrun, 7 stored-block length check.  (Host-callable: tests/png_host_harness.cu runs it on the CPU.)
__host__ __device__ inline int inflate_all(InflateTab& S, const uint8_t* __restrict__ in, int inLen, uint8_t* out, int rawLen)
{
    unsigned long long bb = 0;          // bit buffer (LSB first)
    int bc = 0;                         // valid bits
    const unsigned* ip = (const unsigned*)in;                          // next input word to enter the buffer
    const unsigned* const ipLimit = ip + (inLen + 4) / 4 + 1;          // a refill beyond this word: the stream was overrun
    uint8_t* op = out;                                                 // next output byte
    uint8_t* const oend = out + rawLen;
    int status = 0, last = 0;
    // NEED(n), n <= 32: one aligned word enters the buffer when fewer than n bits are left (bits of the padding behind the stream
    // are never legitimately consumed: checked at the end)
#define NEED(n) if (bc < (n)) { if (ip >= ipLimit) { status = 6; break; } bb |= (unsigned long long)(*ip++) << bc; bc += 32; }
#define CONSUMED() ((int)((const uint8_t*)ip - in) - (bc >> 3))
#define TAKE(var, n) { var = (unsigned)(bb & ((1ull << (n)) - 1)); bb >>= (n); bc -= (n); }
    do {                                                               // (a block for `break` on errors)
        {
            unsigned cmf, flg;
            NEED(16);
            TAKE(cmf, 8); TAKE(flg, 8);
            if ((cmf & 15) != 8 || ((cmf << 8) | flg) % 31 != 0 || (flg & 32)) { status = 1; break; }
        }
        while (!last && !status) {
            unsigned type;
            NEED(3);
            TAKE(last, 1); TAKE(type, 2);
            if (type == 0) {
                const int drop = bc & 7; bb >>= drop; bc -= drop;      // to the byte boundary
                unsigned len, nlen;
                NEED(16);
                TAKE(len, 16);
                NEED(16);
                TAKE(nlen, 16);
                if ((len ^ 0xffffu) != nlen) { status = 7; break; }
                if ((long long)len > oend - op) { status = 5; break; }
                if (CONSUMED() + (int)len > inLen) { status = 6; break; }
                for (unsigned i = 0; i < len; i++) { unsigned v; NEED(8); TAKE(v, 8); *op++ = (uint8_t)v; }
                if (status) break;
                continue;
            }
            if (type == 1) {
                for (int i = 0; i < 144; i++) S.lens[i] = 8;
                for (int i = 144; i < 256; i++) S.lens[i] = 9;
                for (int i = 256; i < 280; i++) S.lens[i] = 7;
                for (int i = 280; i < 288; i++) S.lens[i] = 8;
                huff_build(S.lh, S.lens, 288, S.ltab, PNG_LBITS);
                for (int i = 0; i < 30; i++) S.lens[i] = 5;
                huff_build(S.dh, S.lens, 30, S.dtab, PNG_DBITS);
            } else if (type == 2) {
                unsigned nlen, ndist, ncode;
                NEED(14);
                TAKE(nlen, 5); TAKE(ndist, 5); TAKE(ncode, 4);
                nlen += 257; ndist += 1; ncode += 4;
                if (nlen > 286 || ndist > 30) { status = 3; break; }
                const unsigned char order[19] = { 16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15 };
                for (int i = 0; i < 19; i++) S.lens[i] = 0;
                for (unsigned i = 0; i < ncode; i++) { unsigned v; NEED(3); TAKE(v, 3); S.lens[order[i]] = (unsigned short)v; }
                if (status) break;
                if (huff_build(S.lh, S.lens, 19, S.ltab, 7) < 0) { status = 3; break; }   // the code-length code (7-bit table in ltab)
                unsigned idx = 0;
                unsigned short* tmp = S.lens + 0;                      // the decoded lengths replace the 19 (already consumed by huff_build)
                while (idx < nlen + ndist) {
                    NEED(14);                                          // a code (<= 7 bits) and its extra bits (<= 7)
                    const unsigned short e = S.ltab[bb & 127];
                    if (!e) { status = 3; break; }
                    const int l = e >> 9, sym = e & 511;
                    bb >>= l; bc -= l;
                    if (sym < 16) tmp[idx++] = (unsigned short)sym;
                    else {
                        unsigned rep, val = 0;
                        if (sym == 16) { if (idx == 0) { status = 3; break; } val = tmp[idx - 1]; TAKE(rep, 2); rep += 3; }
                        else if (sym == 17) { TAKE(rep, 3); rep += 3; }
                        else { TAKE(rep, 7); rep += 11; }
                        if (idx + rep > nlen + ndist) { status = 3; break; }
                        while (rep--) tmp[idx++] = (unsigned short)val;
                    }
                }
                if (status) break;
                if (tmp[256] == 0) { status = 3; break; }              // no end-of-block code
                // the distance lengths first (they sit behind the literal/length ones, which building the first code leaves alone)
                if (huff_build(S.dh, tmp + nlen, (int)ndist, S.dtab, PNG_DBITS) < 0) { status = 3; break; }
                if (huff_build(S.lh, tmp, (int)nlen, S.ltab, PNG_LBITS) < 0) { status = 3; break; }
            } else { status = 2; break; }
            // ---- the symbols of a Huffman block ----
            for (;;) {
                NEED(20);                                              // a literal/length code (<= 15 bits) and its extra bits (<= 5)
                const unsigned short e = S.ltab[bb & ((1u << PNG_LBITS) - 1)];
                int l, sym;
                if (e) { l = e >> 9; sym = e & 511; }
                else {                                                 // longer than the table: canonical search, one bit at a time
                    int code = 0, first = 0, index = 0; l = 0; sym = -1;
                    for (int len = 1; len <= 15; len++) {
                        code |= (int)((bb >> (len - 1)) & 1);
                        const int cnt = S.lh.count[len];
                        if (code - cnt < first) { sym = S.lh.symbol[index + (code - first)]; l = len; break; }
                        index += cnt; first += cnt; first <<= 1; code <<= 1;
                    }
                    if (sym < 0) { status = 4; break; }
                }
                bb >>= l; bc -= l;
                if (sym < 256) {
                    if (op >= oend) { status = 5; break; }
                    *op++ = (uint8_t)sym;
                    continue;
                }
                if (sym == 256) break;                                 // end of block
                sym -= 257;
                if (sym >= 29) { status = 4; break; }
                const unsigned short lbase[29] = { 3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258 };
                const unsigned char lext[29] = { 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0 };
                int len = lbase[sym] + (int)(bb & ((1u << lext[sym]) - 1));
                bb >>= lext[sym]; bc -= lext[sym];
                NEED(28);                                              // a distance code (<= 15 bits) and its extra bits (<= 13)
                const unsigned short de = S.dtab[bb & ((1u << PNG_DBITS) - 1)];
                int dl, ds;
                if (de) { dl = de >> 9; ds = de & 511; }
                else {
                    int code = 0, first = 0, index = 0; dl = 0; ds = -1;
                    for (int q = 1; q <= 15; q++) {
                        code |= (int)((bb >> (q - 1)) & 1);
                        const int cnt = S.dh.count[q];
                        if (code - cnt < first) { ds = S.dh.symbol[index + (code - first)]; dl = q; break; }
                        index += cnt; first += cnt; first <<= 1; code <<= 1;
                    }
                    if (ds < 0) { status = 4; break; }
                }
                bb >>= dl; bc -= dl;
                if (ds >= 30) { status = 4; break; }
                const unsigned short dbase[30] = { 1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577 };
                const unsigned char dext[30] = { 0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13 };
                const int dist = dbase[ds] + (int)(bb & ((1u << dext[ds]) - 1));
                bb >>= dext[ds]; bc -= dext[ds];
                if (dist > op - out) { status = 4; break; }
                if (len > oend - op) { status = 5; break; }
                if (dist == 1) {                                       // a run (all that zlib's Z_RLE strategy, OpenCV's default for PNG, ever emits)
                    const uint8_t v = op[-1];
                    for (int i = 0; i < len; i++) op[i] = v;
                } else {
                    const uint8_t* sp_ = op - dist;
                    for (int i = 0; i < len; i++) op[i] = sp_[i];     // byte order matters when dist < len
                }
                op += len;
            }
        }
        if (status) break;
        if (CONSUMED() > inLen) { status = 6; break; }                 // bits beyond the end of the stream were consumed
        if (op != oend) status = 5;
    } while (false);
#undef NEED
#undef TAKE
#undef CONSUMED
    return status;
}

#ifndef PNG_HOST_ONLY
// One single-warp CTA per image; lane 0 decodes (the other lanes only keep the CTA at warp granularity: 32 CTAs, i.e. 32 streams,
// per SM).  Result in img.status (0 ok).
__global__ void __launch_bounds__(32) k_png_inflate(const uint8_t* __restrict__ z, PngImg* imgs, uint8_t* raw, long long rawStride, int W, int H)
{
    __shared__ InflateTab S;
    if (threadIdx.x != 0) return;
    PngImg& im = imgs[blockIdx.x];
    im.status = inflate_all(S, z + im.zOff, (int)im.zLen, raw + (long long)blockIdx.x * rawStride, H * (W * im.bpp + 1));
}

// ---- scanline filters + channel order --------------------------------------------------------------------------------
__device__ __forceinline__ int paeth(int a, int b, int c)
{
    const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}
// One warp per image, rows top-down (Up / Average / Paeth need the row above).  Lane c < bpp owns the chain of channel c along the
// row for Sub / Average / Paeth (each byte depends on the byte bpp positions to its left); None and Up rows are done by all lanes.
// The reconstructed bytes are written back over the raw scanline (so the next row finds its "prior" there), then the row is
// converted to BGR with all lanes.
// The Adler-32 of the inflated bytes is accumulated on the way (a row at a time: A += sum d, B += len A + sum (len - i) d_i) and
// compared with the four bytes that end the zlib stream, as zlib's inflate does: status 8 on a mismatch.
__global__ void __launch_bounds__(32) k_png_unfilter(uint8_t* raw, long long rawStride, PngImg* imgs, const uint8_t* __restrict__ z, int W, int H,
                                                     uint8_t* __restrict__ dst, size_t step, size_t imgStride)
{
    const int lane = threadIdx.x;
    const PngImg im = imgs[blockIdx.x];
    if (im.status) return;
    const int bpp = im.bpp, rowBytes = W * bpp;
    uint8_t* base = raw + (long long)blockIdx.x * rawStride;
    uint8_t* o = dst + (size_t)blockIdx.x * imgStride;
    unsigned adA = 1, adB = 0;                                         // (lane 0's copy is the one that counts)
    for (int y = 0; y < H; y++) {
        uint8_t* cur = base + (long long)y * (rowBytes + 1) + 1;
        const uint8_t* up = y ? cur - (rowBytes + 1) : nullptr;
        {
            const int L = rowBytes + 1;                                // the filter byte belongs to the stream
            unsigned long long sa = 0, sb = 0;
            for (int i = lane; i < L; i += 32) { const unsigned d = cur[i - 1]; sa += d; sb += (unsigned long long)(L - i) * d; }
#pragma unroll
            for (int o2 = 16; o2 > 0; o2 >>= 1) { sa += __shfl_xor_sync(0xffffffffu, sa, o2); sb += __shfl_xor_sync(0xffffffffu, sb, o2); }
            adB = (unsigned)((adB + (unsigned long long)(L % 65521) * adA + sb % 65521ull) % 65521ull);
            adA = (unsigned)((adA + sa) % 65521ull);
        }
        const int ft = cur[-1];
        if (ft == 0) { /* None */ }
        else if (ft == 2) { if (up) for (int i = lane; i < rowBytes; i += 32) cur[i] = (uint8_t)(cur[i] + up[i]); }
        else if (ft == 1) {
            // Sub: a running sum per channel (mod 256) -- every lane sums a stretch of pixels, a warp scan joins the stretches
            const int per = (W + 31) / 32, x0 = min(lane * per, W), x1 = min(x0 + per, W);
            int tot[4] = { 0, 0, 0, 0 };
            for (int x = x0; x < x1; x++)
                for (int ch = 0; ch < bpp; ch++) tot[ch] += cur[x * bpp + ch];
            int offc[4];
            for (int ch = 0; ch < bpp; ch++) {
                int inc = tot[ch];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
                offc[ch] = inc - tot[ch];
            }
            for (int x = x0; x < x1; x++)
                for (int ch = 0; ch < bpp; ch++) { offc[ch] += cur[x * bpp + ch]; cur[x * bpp + ch] = (uint8_t)offc[ch]; }
        }
        else if (lane < bpp) {
            int a = 0, c = 0;                                          // left and upper-left of this channel
            for (int i = lane; i < rowBytes; i += bpp) {
                const int b = up ? up[i] : 0;
                int v = cur[i];
                if (ft == 1) v += a;
                else if (ft == 3) v += (a + b) >> 1;
                else if (ft == 4) v += paeth(a, b, c);
                v &= 255;
                cur[i] = (uint8_t)v;
                a = v; c = b;
            }
        }
        __syncwarp();
        uint8_t* orow = o + (size_t)y * step;
        for (int x = lane; x < W; x += 32) {                            // IMREAD_COLOR: 8UC3 in B, G, R order
            uint8_t r, g, bl;
            if (bpp == 1) r = g = bl = cur[x];
            else { r = cur[x * bpp]; g = cur[x * bpp + 1]; bl = cur[x * bpp + 2]; }
            orow[3 * x] = bl; orow[3 * x + 1] = g; orow[3 * x + 2] = r;
        }
        __syncwarp();
    }
    if (lane == 0) {
        const uint8_t* t = z + im.zOff + im.zLen - 4;                  // big-endian Adler-32 behind the last block
        const unsigned want = ((unsigned)t[0] << 24) | ((unsigned)t[1] << 16) | ((unsigned)t[2] << 8) | t[3];
        if (want != ((adB << 16) | adA)) imgs[blockIdx.x].status = 8;
    }
}

// ---- host ------------------------------------------------------------------------------------------------------------
static unsigned be32(const uint8_t* p) { return ((unsigned)p[0] << 24) | ((unsigned)p[1] << 16) | ((unsigned)p[2] << 8) | p[3]; }

extern "C" int cvr_decode_png_batch(cvr_ctx* c, int n, const uint8_t* const* png, const size_t* png_bytes, int mem_out, uint8_t* bgr, size_t step, size_t image_stride)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    const int W = c->prm.width, H = c->prm.height;
    if (n < 1 || !png || !png_bytes || !bgr || step < (size_t)W * 3 || image_stride < step * (size_t)H || (mem_out != CVR_MEM_HOST && mem_out != CVR_MEM_DEVICE)) {
        c->err = "cvr_decode_png_batch: bad argument"; return CVR_ERR_INVALID_ARG;
    }
    static const uint8_t sig[8] = { 0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a };
    std::vector<PngImg> imgs((size_t)n);
    std::vector<PngSeg> segs;
    long long blobLen = 0, zTotal = 0;
    for (int i = 0; i < n; i++) {
        const uint8_t* f = png[i];
        const size_t len = png_bytes[i];
        if (!f || len < 8 + 25 || memcmp(f, sig, 8) != 0) { c->err = "cvr_decode_png_batch: not a PNG file"; return CVR_ERR_INVALID_ARG; }
        PngImg& im = imgs[(size_t)i];
        im.fileOff = blobLen; im.segFirst = (int)segs.size(); im.segCount = 0; im.zOff = zTotal; im.zLen = 0; im.status = 0; im.colorType = -1; im.bpp = 0;
        size_t p = 8;
        bool end = false;
        while (!end && p + 12 <= len) {
            const unsigned clen = be32(f + p);
            const uint8_t* type = f + p + 4;
            if ((size_t)clen > len - p - 12) { c->err = "cvr_decode_png_batch: truncated PNG chunk"; return CVR_ERR_INVALID_ARG; }
            if (!memcmp(type, "IHDR", 4)) {
                if (clen != 13) { c->err = "cvr_decode_png_batch: bad IHDR"; return CVR_ERR_INVALID_ARG; }
                const uint8_t* d = f + p + 8;
                const unsigned w = be32(d), h = be32(d + 4);
                const int depth = d[8], ct = d[9], interlace = d[12];
                if ((int)w != W || (int)h != H) { c->err = "cvr_decode_png_batch: image size differs from the context's width x height"; return CVR_ERR_INVALID_ARG; }
                if (depth != 8 || !(ct == 0 || ct == 2 || ct == 6) || interlace != 0 || d[10] != 0 || d[11] != 0) {
                    c->err = "cvr_decode_png_batch: only 8-bit non-interlaced gray / RGB / RGBA PNGs are decoded on the device"; return CVR_ERR_UNSUPPORTED;
                }
                im.colorType = ct; im.bpp = ct == 0 ? 1 : (ct == 2 ? 3 : 4);
            } else if (!memcmp(type, "IDAT", 4)) {
                if (clen) { PngSeg s; s.src = blobLen + (long long)p + 8; s.dst = zTotal + im.zLen; s.len = (int)clen; s.img = i; segs.push_back(s); im.zLen += clen; im.segCount++; }
            } else if (!memcmp(type, "IEND", 4)) end = true;
            p += 12 + (size_t)clen;
        }
        if (im.colorType < 0 || im.zLen < 6) { c->err = "cvr_decode_png_batch: PNG without IHDR / IDAT"; return CVR_ERR_INVALID_ARG; }
        blobLen += (long long)len;
        zTotal = (zTotal + im.zLen + 15) & ~15LL;                        // every stream starts 16-byte aligned (the decoder reads whole words)
    }
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    const long long rawStride = ((long long)H * ((long long)W * 4 + 1) + 15) & ~15LL;     // room for RGBA scanlines + filter bytes
    const size_t needBlob = (size_t)blobLen + 16, needZ = (size_t)zTotal + 16, needRaw = (size_t)rawStride * n, needOut = (size_t)W * 3 * H * n;
    auto grow = [&](uint8_t** p, size_t* cap, size_t need) -> int {
        if (*cap >= need) return CVR_OK;
        if (*p) cudaFree(*p);
        *p = nullptr; *cap = 0;
        CVR_CUDA(c, cudaMalloc((void**)p, need + need / 4));
        *cap = need + need / 4;
        return CVR_OK;
    };
    int rc;
    if ((rc = grow(&c->png_blob, &c->png_blob_cap, needBlob)) || (rc = grow(&c->png_z, &c->png_z_cap, needZ)) || (rc = grow(&c->png_raw, &c->png_raw_cap, needRaw)) ||
        (rc = grow(&c->png_tab, &c->png_tab_cap, sizeof(PngImg) * (size_t)n + sizeof(PngSeg) * segs.size() + 64))) return rc;
    uint8_t* dOut = bgr; size_t dStep = step, dStride = image_stride;
    if (mem_out == CVR_MEM_HOST) {
        if ((rc = grow(&c->png_out, &c->png_out_cap, needOut))) return rc;
        dOut = c->png_out; dStep = (size_t)W * 3; dStride = dStep * H;
    }
    // every image inflates to H scanlines of 1 + W * bpp bytes; the images of a batch may differ in colour type
    if ((long long)H * ((long long)W * 4 + 1) > 0x3fffffffLL) { c->err = "cvr_decode_png_batch: image too large for the device decoder"; return CVR_ERR_UNSUPPORTED; }
    cvr_timers_begin(c, 0);
    PngImg* dImgs = (PngImg*)c->png_tab;
    PngSeg* dSegs = (PngSeg*)(c->png_tab + ((sizeof(PngImg) * (size_t)n + 15) & ~(size_t)15));
    CVR_CUDA(c, cudaMemcpyAsync(dImgs, imgs.data(), sizeof(PngImg) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    CVR_CUDA(c, cudaMemcpyAsync(dSegs, segs.data(), sizeof(PngSeg) * segs.size(), cudaMemcpyHostToDevice, c->stream));
    // Large calls run as four chunks of files, each on a stream of its own: the upload of a chunk overlaps the decoding of the
    // chunks before it, and the chunks decode side by side (a stream is latency bound: what counts is how many are in flight).
    const int NCH = n >= 256 ? 4 : 1;
    if (NCH > 1) {
        for (auto& st_ : c->png_stream) if (!st_) CVR_CUDA(c, cudaStreamCreateWithFlags(&st_, cudaStreamNonBlocking));
        for (auto& ev_ : c->png_ev) if (!ev_) CVR_CUDA(c, cudaEventCreateWithFlags(&ev_, cudaEventDisableTiming));
        CVR_CUDA(c, cudaEventRecord(c->png_ev[4], c->stream));          // the tables are on their way
    }
    long long off = 0;
    for (int k = 0; k < NCH; k++) {
        const int i0 = (int)((long long)n * k / NCH), i1 = (int)((long long)n * (k + 1) / NCH), nk = i1 - i0;
        cudaStream_t st = NCH > 1 ? c->png_stream[k] : c->stream;
        if (NCH > 1) CVR_CUDA(c, cudaStreamWaitEvent(st, c->png_ev[4], 0));
        for (int i = i0; i < i1; i++) { CVR_CUDA(c, cudaMemcpyAsync(c->png_blob + off, png[i], png_bytes[i], cudaMemcpyHostToDevice, st)); off += (long long)png_bytes[i]; }
        if (NCH == 1) cvr_stage_mark(c, "png_upload");
        const int s0 = imgs[(size_t)i0].segFirst, s1 = i1 < n ? imgs[(size_t)i1].segFirst : (int)segs.size(), ns = s1 - s0;
        if (ns > 0) {
            k_png_idat<<<(unsigned)(ns < 4096 ? ns : 4096), 128, 0, st>>>(c->png_blob, dSegs + s0, ns, c->png_z);
            LAUNCH_CHECK(c);
        }
        if (NCH == 1) cvr_stage_mark(c, "png_idat");
        k_png_inflate<<<nk, 32, 0, st>>>(c->png_z, dImgs + i0, c->png_raw + (long long)i0 * rawStride, rawStride, W, H);
        LAUNCH_CHECK(c);
        if (NCH == 1) cvr_stage_mark(c, "png_inflate");
        k_png_unfilter<<<nk, 32, 0, st>>>(c->png_raw + (long long)i0 * rawStride, rawStride, dImgs + i0, c->png_z, W, H, dOut + (size_t)i0 * dStride, dStep, dStride);
        LAUNCH_CHECK(c);
        if (NCH == 1) cvr_stage_mark(c, "png_unfilter");
        if (mem_out == CVR_MEM_HOST) {
            if (image_stride == step * (size_t)H) CVR_CUDA(c, cudaMemcpy2DAsync(bgr + (size_t)i0 * image_stride, step, dOut + (size_t)i0 * dStride, dStep, (size_t)W * 3, (size_t)H * nk, cudaMemcpyDeviceToHost, st));
            else for (int i = i0; i < i1; i++) CVR_CUDA(c, cudaMemcpy2DAsync(bgr + (size_t)i * image_stride, step, dOut + (size_t)i * dStride, dStep, (size_t)W * 3, H, cudaMemcpyDeviceToHost, st));
        }
        if (NCH > 1) {
            CVR_CUDA(c, cudaEventRecord(c->png_ev[k], st));
            CVR_CUDA(c, cudaStreamWaitEvent(c->stream, c->png_ev[k], 0));
        }
    }
    if (NCH > 1) cvr_stage_mark(c, "png_pipeline");
    CVR_CUDA(c, cudaMemcpyAsync(imgs.data(), dImgs, sizeof(PngImg) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) { c->err = std::string("stream sync: ") + cudaGetErrorString(e); return CVR_ERR_CUDA; }
    cvr_timers_end(c, 0);
    for (int i = 0; i < n; i++)
        if (imgs[(size_t)i].status) { c->err = "cvr_decode_png_batch: corrupt or unsupported deflate stream in image " + std::to_string(i) + " (code " + std::to_string(imgs[(size_t)i].status) + ")"; return CVR_ERR_INVALID_ARG; }
    return CVR_OK;
}
#endif  // PNG_HOST_ONLY
