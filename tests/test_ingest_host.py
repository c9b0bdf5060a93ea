"""CPU-only tests of the mask file decoder behind vcb_load_mask_file / vcb_carve_mask_files
(csrc/ingest.cu): 8-bit grey PNG and binary PGM.  The PNG fixtures are written here with Python's
zlib - every scanline filter of the PNG specification, stored / fixed-Huffman / dynamic-Huffman
DEFLATE blocks, several IDAT chunks - so the expected pixels are known exactly; the reference's
own loader (Image1b::Load, image.h:76-101, stb_image) returns the same bytes for this format."""
import os
import struct
import zlib

import numpy as np
import pytest

from vacancy_b200 import capi, load_mask_file


def _chunk(kind, body):
    return struct.pack(">I", len(body)) + kind + body + struct.pack(">I", zlib.crc32(kind + body))


def _paeth(a, b, c):
    p = a + b - c
    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
    return a if pa <= pb and pa <= pc else (b if pb <= pc else c)


def filter_rows(img, filters):
    """PNG scanline filtering (specification section 9) of a uint8 [H, W] image, bpp = 1."""
    h, w = img.shape
    out = bytearray()
    for y in range(h):
        f = filters[y % len(filters)]
        row = img[y].astype(np.int32)
        left = np.concatenate([[0], row[:-1]])
        up = img[y - 1].astype(np.int32) if y else np.zeros(w, np.int32)
        ul = np.concatenate([[0], up[:-1]])
        if f == 0:
            r = row
        elif f == 1:
            r = row - left
        elif f == 2:
            r = row - up
        elif f == 3:
            r = row - ((left + up) >> 1)
        else:
            r = row - np.array([_paeth(int(a), int(b), int(c)) for a, b, c in zip(left, up, ul)])
        out += bytes([f]) + (r & 255).astype(np.uint8).tobytes()
    return bytes(out)


def write_png(path, img, filters=(0,), level=6, strategy=zlib.Z_DEFAULT_STRATEGY, idat_split=1,
              color_type=0, depth=8, interlace=0):
    h, w = img.shape[:2]
    co = zlib.compressobj(level, zlib.DEFLATED, 15, 9, strategy)
    data = co.compress(filter_rows(img, filters)) + co.flush()
    parts = [data[i * len(data) // idat_split:(i + 1) * len(data) // idat_split]
             for i in range(idat_split)]
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n")
        f.write(_chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, color_type, 0, 0, interlace)))
        f.write(_chunk(b"tEXt", b"Comment\x00made by the test"))
        for p in parts:
            f.write(_chunk(b"IDAT", p))
        f.write(_chunk(b"IEND", b""))


def test_png_every_filter_and_block_type(tmp_path):
    rng = np.random.default_rng(3)
    yy, xx = np.mgrid[0:97, 0:133]
    blob = (((yy - 50) ** 2 + (xx - 60) ** 2 < 900) * 255).astype(np.uint8)   # a silhouette
    noise = rng.integers(0, 256, (61, 75), dtype=np.uint8)                     # incompressible
    grad = ((yy * 2 + xx * 3) % 256).astype(np.uint8)
    cases = [
        (blob, (0,), 6, zlib.Z_DEFAULT_STRATEGY, 1),
        (blob, (1, 2, 3, 4, 0), 9, zlib.Z_DEFAULT_STRATEGY, 3),
        (grad, (4,), 6, zlib.Z_DEFAULT_STRATEGY, 1),
        (grad, (3, 1), 1, zlib.Z_FIXED, 2),          # fixed Huffman blocks
        (noise, (0, 4), 0, zlib.Z_DEFAULT_STRATEGY, 1),  # stored blocks
        (noise, (2,), 6, zlib.Z_HUFFMAN_ONLY, 1),
        (blob[:1, :], (1,), 6, zlib.Z_DEFAULT_STRATEGY, 1),
        (blob[:, :1].copy(), (4,), 6, zlib.Z_DEFAULT_STRATEGY, 1),
    ]
    for i, (img, filters, level, strategy, split) in enumerate(cases):
        p = tmp_path / ("m%d.png" % i)
        write_png(p, np.ascontiguousarray(img), filters, level, strategy, split)
        got = load_mask_file(p)
        assert got.shape == img.shape and np.array_equal(got, img), i


def test_large_mask_png_and_pgm(tmp_path):
    """A 2048 x 2048 silhouette (the benchmark's image size): long matches, many dynamic blocks."""
    from vacancy_b200 import synthetic as S
    mask = S.silhouette(S.circle_cameras(4, 2048)[1], S.DUMBBELL)
    write_png(tmp_path / "big.png", mask, (0, 2), 6)
    assert np.array_equal(load_mask_file(tmp_path / "big.png"), mask)
    with open(tmp_path / "big.pgm", "wb") as f:
        f.write(b"P5\n# a comment line\n2048 2048\n255\n" + mask.tobytes())
    assert np.array_equal(load_mask_file(tmp_path / "big.pgm"), mask)


def test_refused_files(tmp_path):
    lib = capi.load_library()
    img = np.zeros((8, 8), np.uint8)
    write_png(tmp_path / "rgb.png", np.zeros((8, 24), np.uint8), color_type=2)
    write_png(tmp_path / "deep.png", np.zeros((8, 16), np.uint8), depth=16)
    write_png(tmp_path / "adam7.png", img, interlace=1)
    for name in ("rgb.png", "deep.png", "adam7.png"):
        with pytest.raises(capi.VcbError) as e:
            load_mask_file(tmp_path / name)
        assert e.value.code == -4 and "8-bit greyscale" in str(e.value)
    write_png(tmp_path / "ok.png", img)
    raw = open(tmp_path / "ok.png", "rb").read()
    open(tmp_path / "cut.png", "wb").write(raw[:len(raw) - 20])
    open(tmp_path / "text.txt", "w").write("not an image")
    open(tmp_path / "p2.pgm", "w").write("P5\n4 4\n65535\n")
    for name, code in (("cut.png", -1), ("text.txt", -4), ("p2.pgm", -4), ("missing.png", -1)):
        with pytest.raises(capi.VcbError) as e:
            load_mask_file(tmp_path / name)
        assert e.value.code == code, name
    assert lib.vcb_load_mask_file(None, None, 0, None, None) == -1


def test_bundled_bunny_masks_when_present(bunny):
    """The reference's own data/mask_*.png (copied next to the compiled reference, git-ignored):
    the decoder must give the pixels the reference's stb loader gave (tests/golden/bunny_inputs.npz)."""
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref",
                        "data")
    if not os.path.exists(os.path.join(root, "mask_00000.png")):
        pytest.skip("oracle/_ref/data not present")
    for i in range(6):
        assert np.array_equal(load_mask_file(os.path.join(root, "mask_%05d.png" % i)), bunny.masks[i])


def test_damaged_files_never_crash(tmp_path):
    """The decoder reads files it did not write: random byte damage anywhere in a PNG (header, zlib
    stream, Huffman tables, chunk lengths) must end in a clean error or in some image, never in a
    crash or an out-of-bounds write (the output buffer is sized from the damaged header itself)."""
    rng = np.random.default_rng(99)
    yy, xx = np.mgrid[0:120, 0:160]
    img = (((yy - 60) ** 2 + (xx - 70) ** 2 < 1500) * 255).astype(np.uint8)
    img[::7, ::5] = rng.integers(0, 256, img[::7, ::5].shape, dtype=np.uint8)
    write_png(tmp_path / "good.png", img, (0, 1, 2, 3, 4), 6, idat_split=2)
    good = open(tmp_path / "good.png", "rb").read()
    outcomes = {"ok": 0, "error": 0}
    for t in range(400):
        data = bytearray(good)
        for _ in range(int(rng.integers(1, 4))):
            data[int(rng.integers(8, len(data)))] = int(rng.integers(0, 256))
        if t % 10 == 0:
            data = data[:int(rng.integers(20, len(data)))]
        open(tmp_path / "bad.png", "wb").write(bytes(data))
        try:
            out = load_mask_file(tmp_path / "bad.png")
            assert out.ndim == 2
            outcomes["ok"] += 1
        except capi.VcbError as e:
            assert e.code in (-1, -4)
            outcomes["error"] += 1
    assert outcomes["error"] > 50  # most damage is detected (sizes, codes, stream length)
