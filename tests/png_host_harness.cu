// Test infrastructure: runs the sequential inflate routine of master-thesis_b200/csrc/k_png.cu (inflate_run, a __host__ __device__
// function) on the CPU, driving its input ring / output window exactly as the kernel's cooperative loop does, so that the bit-stream
// logic is covered by the CPU test suite (tests/test_png_host.py compares it with zlib).  Built with nvcc -DPNG_HOST_ONLY.
#define PNG_HOST_ONLY 1
#include "../master-thesis_b200/csrc/k_png.cu"

extern "C" int png_inflate_host(const uint8_t* z, long long zlen, uint8_t* out, long long rawLen)
{
    InflateSm* S = new InflateSm;
    for (int i = 0; i < PNG_INRING; i++) S->in[i] = i < zlen ? z[i] : 0;
    InflateState st;
    memset(&st, 0, sizeof(st));
    int inBase = 0, flushed = 0;
    int result;
    for (;;) {
        if (!st.done && !st.status) inflate_run(st, *S, inBase + PNG_INRING / 2, flushed + PNG_WIN / 2, (int)zlen, (int)rawLen);
        const int consumed = st.b.pos - (st.b.bc >> 3);
        const int upto = (st.done || st.status) ? st.outPos : (st.outPos / (PNG_WIN / 2)) * (PNG_WIN / 2);
        for (int i = flushed; i < upto && i < rawLen; i++) out[i] = S->win[i & (PNG_WIN - 1)];
        if (upto > flushed) flushed = upto;
        if (st.done || st.status) { result = st.status ? st.status : (st.outPos == rawLen ? 0 : 5); break; }
        while (consumed >= inBase + PNG_INRING / 2) {
            const int src = inBase + PNG_INRING;
            for (int i = 0; i < PNG_INRING / 2; i++) S->in[(src + i) & (PNG_INRING - 1)] = (src + i) < zlen ? z[src + i] : 0;
            inBase += PNG_INRING / 2;
        }
    }
    delete S;
    return result;
}
