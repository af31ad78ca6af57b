// Test infrastructure: runs the inflate routine of master-thesis_b200/csrc/k_png.cu (inflate_all, a __host__ __device__ function:
// the very code lane 0 of the kernel executes) on the CPU, so that the bit-stream logic is covered by the CPU test suite
// (tests/test_png_host.py compares it with zlib).  Built with nvcc -DPNG_HOST_ONLY.
#define PNG_HOST_ONLY 1
#include "../master-thesis_b200/csrc/k_png.cu"

extern "C" int png_inflate_host(const uint8_t* z, long long zlen, uint8_t* out, long long rawLen)
{
    InflateTab* S = new InflateTab;
    // the decoder reads its input a 32-bit word at a time: 4-byte aligned, readable 8 bytes past the end (as cvr_decode_png_batch lays it out)
    unsigned* buf = new unsigned[(size_t)(zlen + 16) / 4 + 1];
    memset(buf, 0, ((size_t)(zlen + 16) / 4 + 1) * 4);
    memcpy(buf, z, (size_t)zlen);
    uint8_t* o = new uint8_t[(size_t)rawLen + 1];
    const int result = inflate_all(*S, (const uint8_t*)buf, (int)zlen, o, (int)rawLen);
    memcpy(out, o, (size_t)rawLen);
    delete[] o; delete[] buf; delete S;
    return result;
}
