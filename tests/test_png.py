"""GPU: cvr_decode_png_batch -- the cv::imdecode(png, IMREAD_COLOR) step of MessageConverter::ConvertStereoMessage
(MessageConverter.cpp:65-66) on the device.  PNG is lossless: the decoded frames must equal cv2.imdecode's byte for byte.
Files come from cv2.imencode (what the reference's streamers call) at several zlib levels, and from a hand-rolled PNG writer
that forces what OpenCV's encoder never emits: every scanline filter, fixed-Huffman and stored blocks, long-distance matches,
IDAT chunks cut at odd places."""
import struct
import zlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
cv2 = pytest.importorskip("cv2")

H, W = 96, 400


def _frames(oracle, n, smooth=False):
    out = []
    for i in range(n):
        l, _ = oracle.synth_pair(H, W, 64, seed=600 + i, color=True)
        if smooth:                                              # large flat areas: long matches, few literals
            l = (l // 32 * 32).astype(np.uint8)
            l[H // 3:, W // 4:] = l[H // 3, W // 4]
        out.append(l)
    return out


def _chunk(tag, data):
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xffffffff)


def write_png(img_rgb, filters, level=9, strategy=zlib.Z_DEFAULT_STRATEGY, idat=8192, color_type=None):
    """A PNG writer under the test's control: img_rgb is (H, W, C) with C in {1, 3, 4} in PNG channel order."""
    h, w, ch = img_rgb.shape
    ct = color_type if color_type is not None else {1: 0, 3: 2, 4: 6}[ch]
    raw = bytearray()
    prev = np.zeros(w * ch, np.int32)
    for y in range(h):
        cur = img_rgb[y].reshape(-1).astype(np.int32)
        ft = filters[y % len(filters)]
        left = np.concatenate([np.zeros(ch, np.int32), cur[:-ch]])
        ul = np.concatenate([np.zeros(ch, np.int32), prev[:-ch]])
        if ft == 0:
            f = cur
        elif ft == 1:
            f = cur - left
        elif ft == 2:
            f = cur - prev
        elif ft == 3:
            f = cur - ((left + prev) >> 1)
        else:
            p = left + prev - ul
            pa, pb, pc = np.abs(p - left), np.abs(p - prev), np.abs(p - ul)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, ul))
            f = cur - pred
        raw.append(ft)
        raw += (f & 255).astype(np.uint8).tobytes()
        prev = cur
    co = zlib.compressobj(level, zlib.DEFLATED, 15, 9, strategy)
    z = co.compress(bytes(raw)) + co.flush()
    out = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, ct, 0, 0, 0))
    for o in range(0, len(z), idat):
        out += _chunk(b"IDAT", z[o:o + idat])
    return out + _chunk(b"tEXt", b"k\x00v") + _chunk(b"IEND", b"")


@pytest.mark.parametrize("level", [0, 1, 3, 9])
def test_opencv_encoded_frames_decode_bit_exact(cvr, oracle, level):
    rec = cvr.Reconstruct3D(W, H, filter_only=True, max_points=16)
    imgs = _frames(oracle, 5)
    files = [cv2.imencode(".png", im, [cv2.IMWRITE_PNG_COMPRESSION, level])[1].tobytes() for im in imgs]
    got = rec.DecodePNG(files)
    for i, f in enumerate(files):
        want = cv2.imdecode(np.frombuffer(f, np.uint8), cv2.IMREAD_COLOR)
        assert np.array_equal(want, imgs[i])                     # (lossless round trip on the host side)
        assert np.array_equal(got[i], want), "image %d, level %d" % (i, level)
    rec.close()


def test_gray_and_rgba_follow_imread_color(cvr, oracle):
    rec = cvr.Reconstruct3D(W, H, filter_only=True, max_points=16)
    bgr = _frames(oracle, 1)[0]
    gray = cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)
    bgra = np.dstack([bgr, np.full((H, W), 77, np.uint8)])
    files = [cv2.imencode(".png", gray)[1].tobytes(), cv2.imencode(".png", bgra)[1].tobytes(), cv2.imencode(".png", bgr)[1].tobytes()]
    got = rec.DecodePNG(files)                                  # a batch that mixes colour types
    for i, f in enumerate(files):
        assert np.array_equal(got[i], cv2.imdecode(np.frombuffer(f, np.uint8), cv2.IMREAD_COLOR)), i
    rec.close()


@pytest.mark.parametrize("filters,level,strategy,idat", [
    ([0], 9, zlib.Z_DEFAULT_STRATEGY, 8192), ([1], 6, zlib.Z_DEFAULT_STRATEGY, 8192), ([2], 6, zlib.Z_DEFAULT_STRATEGY, 8192),
    ([3], 6, zlib.Z_DEFAULT_STRATEGY, 8192), ([4], 6, zlib.Z_DEFAULT_STRATEGY, 8192), ([4, 3, 2, 1, 0], 9, zlib.Z_DEFAULT_STRATEGY, 1000),
    ([1, 4], 9, zlib.Z_FIXED, 4099), ([0, 2], 9, zlib.Z_HUFFMAN_ONLY, 7), ([3, 4, 1], 1, zlib.Z_RLE, 333), ([2, 4], 0, zlib.Z_DEFAULT_STRATEGY, 65536)])
def test_every_filter_and_block_type(cvr, oracle, filters, level, strategy, idat):
    rec = cvr.Reconstruct3D(W, H, filter_only=True, max_points=16)
    files, wants = [], []
    for k, im in enumerate(_frames(oracle, 3, smooth=True)):
        rgb = im[:, :, ::-1]
        src = [rgb, rgb[:, :, :1], np.dstack([rgb, np.full((H, W), 9, np.uint8)])][k]
        f = write_png(np.ascontiguousarray(src), filters, level, strategy, idat)
        want = cv2.imdecode(np.frombuffer(f, np.uint8), cv2.IMREAD_COLOR)
        assert want is not None and want.shape == (H, W, 3)
        files.append(f); wants.append(want)
    got = rec.DecodePNG(files)
    for i in range(3):
        assert np.array_equal(got[i], wants[i]), i
    rec.close()


def test_large_call_runs_as_chunks_on_several_streams(cvr, oracle):
    """From 256 files per call on, the files are uploaded and decoded in four chunks on streams of their own: same frames, in
    the caller's order, to host and to device memory; a damaged file in the last chunk is still reported."""
    import torch
    rec = cvr.Reconstruct3D(W, H, filter_only=True, max_points=16)
    base = _frames(oracle, 7)
    gray = cv2.cvtColor(base[0], cv2.COLOR_BGR2GRAY)
    enc = [cv2.imencode(".png", im, [cv2.IMWRITE_PNG_COMPRESSION, 1 + i % 3])[1].tobytes() for i, im in enumerate(base)]
    enc.append(cv2.imencode(".png", gray)[1].tobytes())               # a batch that mixes colour types, odd sizes per chunk
    wants = [cv2.imdecode(np.frombuffer(f, np.uint8), cv2.IMREAD_COLOR) for f in enc]
    n = 263
    order = [(5 * i + i // 8) % len(enc) for i in range(n)]
    files = [enc[k] for k in order]
    got = rec.DecodePNG(files)
    for i, k in enumerate(order):
        assert np.array_equal(got[i], wants[k]), i
    dev = torch.zeros((n, H, W, 3), dtype=torch.uint8, device="cuda")
    rec.DecodePNG(files, device_out=dev.data_ptr())
    assert np.array_equal(dev.cpu().numpy(), got)
    bad = bytearray(files[-2]); bad[bad.index(b"IDAT") + 4] ^= 0x07       # the zlib header of the stream (chunk CRCs are not checked on the device)
    with pytest.raises(cvr.CvrError):
        rec.DecodePNG(files[:-2] + [bytes(bad), files[-1]])
    # damage inside the compressed data that may well still inflate to the right length: caught by the stream's Adler-32
    for pos in (100, 1000, 5000):
        bad = bytearray(files[3]); bad[bad.index(b"IDAT") + 4 + pos] ^= 0x10
        with pytest.raises(cvr.CvrError):
            rec.DecodePNG(files[:3] + [bytes(bad)] + files[4:9])
    got2 = rec.DecodePNG(files[:300 - n] if n > 300 else files[:257])  # the decoder is usable after a failed call
    assert np.array_equal(got2, got[:len(got2)])
    rec.close()


def test_refuses_what_it_does_not_decode(cvr, oracle):
    rec = cvr.Reconstruct3D(W, H, filter_only=True, max_points=16)
    im = _frames(oracle, 1)[0]
    good = cv2.imencode(".png", im)[1].tobytes()
    with pytest.raises(cvr.CvrError) as e:
        rec.DecodePNG([cv2.imencode(".png", im.astype(np.uint16) * 257)[1].tobytes()])      # 16-bit samples
    assert e.value.code == -2
    with pytest.raises(cvr.CvrError) as e:
        rec.DecodePNG([cv2.imencode(".png", im[:, :W - 2])[1].tobytes()])                   # another size
    assert e.value.code == -1
    with pytest.raises(cvr.CvrError):
        rec.DecodePNG([good[:len(good) // 2]])                                             # cut in the middle of a chunk
    bad = bytearray(good)
    k = good.index(b"IDAT") + 4 + 40
    bad[k:k + 64] = bytes(64)                                                              # a damaged deflate stream must not hang or crash
    try:
        rec.DecodePNG([bytes(bad)])
    except cvr.CvrError:
        pass
    assert np.array_equal(rec.DecodePNG([good])[0], im)                                    # and the context still works afterwards
    rec.close()


def test_decoded_frames_feed_the_path_without_touching_the_host(cvr, oracle):
    """PNG files -> device frames -> cvr_submit_batch(CVR_MEM_DEVICE): same clouds as from the decoded host arrays."""
    import torch
    D, bs, B = 64, 9, 3
    pairs = [oracle.synth_pair(H, W, D, seed=700 + i, color=True) for i in range(B)]
    L = np.stack([p[0] for p in pairs]); R = np.stack([p[1] for p in pairs])
    cfg = dict(block_size=bs, num_disparities=D, uniqueness_ratio=10, speckle_range=2, speckle_window_size=100, pre_filter_cap=10,
               min_disparity=0, outlier_min_k=8, outlier_std_threshold=0.5)
    rec = cvr.Reconstruct3D(W, H, config=cfg, max_batch=B, voxel_leaf=0.05)
    _, want = rec.ProcessFrames(L, R)
    dev = torch.device("cuda", 0)
    dL = torch.empty((B, H, W, 3), dtype=torch.uint8, device=dev); dR = torch.empty_like(dL)
    rec.DecodePNG([cv2.imencode(".png", im)[1].tobytes() for im in L], device_out=dL.data_ptr())
    rec.DecodePNG([cv2.imencode(".png", im)[1].tobytes() for im in R], device_out=dR.data_ptr())
    assert np.array_equal(dL.cpu().numpy(), L) and np.array_equal(dR.cpu().numpy(), R)
    cap = W * H * B
    packed = torch.zeros((cap, 16), dtype=torch.uint8, device=dev); off = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    io = cvr.Batch()
    io.mem, io.channels, io.stages = cvr.MEM_DEVICE, 3, 15
    io.left, io.right = dL.data_ptr(), dR.data_ptr()
    io.left_step = io.right_step = W * 3
    io.left_frame_stride = io.right_frame_stride = W * H * 3
    io.cloud_packed, io.cloud_packed_capacity, io.cloud_offsets = packed.data_ptr(), cap, off.data_ptr()
    rec.submit_batch_raw(io, B); rec.wait_batch()
    o = off.cpu().numpy(); flat = packed.cpu().numpy().reshape(-1).view(cvr.POINT_DTYPE)
    for i in range(B):
        assert np.array_equal(flat[o[i]:o[i + 1]], want[i])
    rec.close()
