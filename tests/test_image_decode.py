"""Image decoding for LoadTexture: host/image.cpp against the reference's own vendored stb_image.h
(stbi_loadf(path, &w, &h, &n, 4), scene.cpp:294-297) compiled into oracle/_ref/libref_host.so - bit for bit on the float
pixels, for PNG (every colour type / depth, tRNS, Adam7), JPEG (baseline, progressive, subsampling, restart intervals,
greyscale) and Radiance HDR (run-length and flat).  Files are written here with Pillow or by hand.  CPU only."""
import struct
import zlib

import numpy as np
import pytest

from oracle.ref import host_binding as H
from path_tracer_b200 import scene as S

PIL = pytest.importorskip("PIL.Image")
pytestmark = pytest.mark.skipif(not H.available(), reason="oracle/_ref/libref_host.so cannot be built here")


def picture(w, h, seed=0):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    base = np.stack([128 + 100 * np.sin(x / 7.0), 128 + 100 * np.cos(y / 5.0), (x * 3 + y * 5) % 256, 255 * ((x // 4 + y // 4) % 2)], -1)
    return np.clip(base + rng.integers(-20, 20, size=(h, w, 4)), 0, 255).astype(np.uint8)


def same(path):
    ours, ref = S.decode_image(str(path)), H.stbi_loadf(str(path))
    assert ref is not None, "the reference's stb_image cannot read %s" % path
    assert ours is not None, "host/image.cpp cannot read %s" % path
    assert ours.shape == ref.shape
    if ours.tobytes() != ref.tobytes():
        d = np.argwhere((ours.view(np.uint32) != ref.view(np.uint32)).any(axis=-1))
        raise AssertionError("%s: %d of %d pixels differ, first at %s: %s vs %s" % (path, len(d), ours.shape[0] * ours.shape[1], d[0], ours[tuple(d[0])], ref[tuple(d[0])]))
    return ours


@pytest.mark.parametrize("mode", ["L", "LA", "RGB", "RGBA", "P", "P-trns", "1", "I;16", "RGB-trns", "L-trns"])
@pytest.mark.parametrize("size", [(37, 23), (1, 1), (8, 300)])
def test_png_matches_stb_image(tmp_path, mode, size):
    w, h = size
    a = picture(w, h, 1)
    f = tmp_path / "t.png"
    if mode == "P":
        PIL.fromarray(a[..., :3]).quantize(17).save(f)
    elif mode == "P-trns":
        im = PIL.fromarray(a[..., :3]).quantize(5)
        im.save(f, transparency=bytes([0, 128, 255, 7, 200]))
    elif mode == "1":
        PIL.fromarray(a[..., 0] > 128).save(f)
    elif mode == "I;16":
        PIL.fromarray((a[..., 0].astype(np.uint16) * 257 + a[..., 1]) & 0xFFFF).save(f)
    elif mode == "RGB-trns":
        b = a[..., :3].copy()
        b[::3, ::2] = (10, 20, 30)
        PIL.fromarray(b).save(f, transparency=(10, 20, 30))
    elif mode == "L-trns":
        b = a[..., 0].copy()
        b[::2, ::3] = 77
        PIL.fromarray(b).save(f, transparency=77)
    else:
        PIL.fromarray(a if mode == "RGBA" else a[..., :3] if mode == "RGB" else a[..., 0] if mode == "L" else a[..., [0, 3]], mode).save(f)
    px = same(f)
    assert px.shape == (h, w, 4)


def png_chunk(tag, data):
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)


def write_png_raw(path, w, h, depth, ctype, rows_of, interlace, extra=b"", idat_split=1):
    """rows_of(pw, ph, xs, ys) -> list of raw scanlines (bytes, no filter byte); a random filter type is applied here."""
    rng = np.random.default_rng(w * 131 + h)
    channels = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    bpp = max(1, channels * depth // 8)

    def filt(rows):
        out, prev = b"", None
        for r in rows:
            r = bytearray(r)
            t = int(rng.integers(0, 5))
            p = prev if prev is not None else bytearray(len(r))
            f = bytearray(len(r))
            for i in range(len(r)):
                a = r[i - bpp] if i >= bpp else 0
                b = p[i]
                c = p[i - bpp] if i >= bpp else 0
                if t == 0:
                    pr = 0
                elif t == 1:
                    pr = a
                elif t == 2:
                    pr = b
                elif t == 3:
                    pr = (a + b) >> 1
                else:
                    pa, pb, pc = abs(b - c), abs(a - c), abs(a + b - 2 * c)
                    pr = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                f[i] = (r[i] - pr) & 255
            out += bytes([t]) + bytes(f)
            prev = r
        return out
    raw = b""
    if not interlace:
        raw = filt(rows_of(w, h, lambda i: i, lambda j: j))
    else:
        for xo, yo, xs, ys in ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)):
            pw, ph = (w - xo + xs - 1) // xs, (h - yo + ys - 1) // ys
            if pw > 0 and ph > 0:
                raw += filt(rows_of(pw, ph, lambda i, xo=xo, xs=xs: xo + i * xs, lambda j, yo=yo, ys=ys: yo + j * ys))
    z = zlib.compress(raw, 6)
    cut = [len(z) * k // idat_split for k in range(idat_split + 1)]
    with open(path, "wb") as o:
        o.write(b"\x89PNG\r\n\x1a\n" + png_chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 1 if interlace else 0)) + extra)
        for k in range(idat_split):
            o.write(png_chunk(b"IDAT", z[cut[k]:cut[k + 1]]))
        o.write(png_chunk(b"IEND", b""))


@pytest.mark.parametrize("interlace", [False, True])
@pytest.mark.parametrize("depth,ctype", [(1, 0), (2, 0), (4, 0), (8, 0), (16, 0), (8, 2), (16, 2), (1, 3), (2, 3), (4, 3), (8, 3), (8, 4), (16, 4), (8, 6), (16, 6)])
def test_png_every_type_and_depth_with_filters_and_adam7(tmp_path, depth, ctype, interlace):
    w, h = 21, 13
    rng = np.random.default_rng(depth * 10 + ctype)
    channels = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    samples = rng.integers(0, 1 << depth, size=(h, w, channels))

    def rows_of(pw, ph, xs, ys):
        rows = []
        for j in range(ph):
            vals = [int(samples[ys(j), xs(i), c]) for i in range(pw) for c in range(channels)]
            if depth == 16:
                rows.append(b"".join(struct.pack(">H", v) for v in vals))
            elif depth == 8:
                rows.append(bytes(vals))
            else:
                bits = "".join(format(v, "0%db" % depth) for v in vals)
                bits += "0" * (-len(bits) % 8)
                rows.append(bytes(int(bits[k:k + 8], 2) for k in range(0, len(bits), 8)))
        return rows
    extra = b""
    if ctype == 3:
        n = 1 << depth
        extra = png_chunk(b"PLTE", bytes(rng.integers(0, 256, size=3 * n).tolist())) + png_chunk(b"tRNS", bytes(rng.integers(0, 256, size=n // 2 + 1).tolist()))
    elif ctype == 0:
        extra = png_chunk(b"tRNS", struct.pack(">H", int(samples[0, 0, 0])))
    elif ctype == 2:
        extra = png_chunk(b"tRNS", struct.pack(">HHH", *[int(v) for v in samples[1, 2]]))
    f = tmp_path / "t.png"
    write_png_raw(f, w, h, depth, ctype, rows_of, interlace, extra, idat_split=3)
    px = same(f)
    if ctype in (0, 2):
        assert px[..., 3].min() == 0.0                      # the colour key made something transparent


@pytest.mark.parametrize("kind", ["444", "420", "422", "grey", "progressive", "progressive-420", "restart", "odd-size", "q30"])
def test_jpeg_matches_stb_image(tmp_path, kind):
    w, h = (67, 45) if kind == "odd-size" else (96, 64)
    a = picture(w, h, 3)[..., :3]
    f = tmp_path / "t.jpg"
    im = PIL.fromarray(a)
    kw = dict(quality=30 if kind == "q30" else 88)
    if kind == "grey":
        im = im.convert("L")
    elif kind in ("444", "restart", "progressive"):
        kw["subsampling"] = 0
    elif kind == "422":
        kw["subsampling"] = 1
    else:
        kw["subsampling"] = 2
    if kind.startswith("progressive"):
        kw["progressive"] = True
    if kind == "restart":
        kw["restart_marker_blocks"] = 5
    try:
        im.save(f, "JPEG", **kw)
    except TypeError:
        kw.pop("restart_marker_blocks", None)
        im.save(f, "JPEG", **kw)
    px = same(f)
    assert px.shape == (h, w, 4) and (px[..., 3] == 1.0).all()


def write_hdr(path, rgbe, rle, magic=b"#?RADIANCE"):
    h, w = rgbe.shape[:2]
    with open(path, "wb") as o:
        o.write(magic + b"\n# made by a test\nFORMAT=32-bit_rle_rgbe\nEXPOSURE=1.0\n\n-Y %d +X %d\n" % (h, w))
        for j in range(h):
            if not rle or w < 8:                            # scanlines shorter than 8 pixels are never run-length coded
                o.write(rgbe[j].tobytes())
                continue
            o.write(bytes([2, 2, w >> 8, w & 255]))
            for c in range(4):
                row = rgbe[j, :, c]
                i = 0
                while i < w:
                    run = 1
                    while i + run < w and run < 127 and row[i + run] == row[i]:
                        run += 1
                    if run >= 3:
                        o.write(bytes([128 + run, int(row[i])]))
                        i += run
                    else:
                        k = i
                        while k < w and k - i < 128 and not (k + 2 < w and row[k] == row[k + 1] == row[k + 2]):
                            k += 1
                        k = max(k, i + 1)
                        o.write(bytes([k - i]) + row[i:k].tobytes())
                        i = k


@pytest.mark.parametrize("w,h,rle,magic", [(40, 17, True, b"#?RADIANCE"), (40, 17, False, b"#?RADIANCE"), (5, 9, True, b"#?RGBE"), (300, 4, True, b"#?RADIANCE")])
def test_radiance_hdr_matches_stb_image(tmp_path, w, h, rle, magic):
    rng = np.random.default_rng(w)
    rgbe = rng.integers(0, 256, size=(h, w, 4)).astype(np.uint8)
    rgbe[..., 3] = rng.integers(100, 150, size=(h, w))
    rgbe[:, : w // 3, 0] = 200                              # runs for the run-length coder
    rgbe[0, 0, 3] = 0                                       # exponent 0: black
    f = tmp_path / "t.hdr"
    write_hdr(f, rgbe, rle, magic)
    px = same(f)
    assert px[0, 0, :3].max() == 0.0 and (px[..., 3] == 1.0).all() and px.max() > 100.0


def test_load_texture_and_mtl_maps(tmp_path):
    """LoadTexture / LoadModelAsPrefab take PNG, JPEG and HDR files: a textured OBJ and an HDR sky end up in the atlas."""
    S.ensure_spectrum_table()
    a = picture(32, 16, 5)
    PIL.fromarray(a[..., :3]).save(tmp_path / "kd.png")
    rgbe = np.full((8, 16, 4), 128, np.uint8)
    rgbe[..., 0] = np.arange(16, dtype=np.uint8) * 10
    write_hdr(tmp_path / "sky.hdr", rgbe, True)
    with open(tmp_path / "m.mtl", "w") as o:
        o.write("newmtl tex\nKd 1 1 1\nmap_Kd kd.png\n")
    with open(tmp_path / "m.obj", "w") as o:
        o.write("mtllib m.mtl\no quad\nv 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvt 0 0\nvt 1 0\nvt 1 1\nvt 0 1\nusemtl tex\nf 1/1 2/2 3/3\nf 1/1 3/3 4/4\n")
    sc = S.Scene()
    sc.set_atlas_size(64, 64)
    prefab = sc.load_model(str(tmp_path / "m.obj"))
    sc.create_prefab_instance(prefab)
    sky = sc.load_texture(str(tmp_path / "sky.hdr"), S.TEXTURE_RADIANCE)
    sc.set_root(sky_texture=sky)
    sc.create_camera((0, -3, 1), (0, 0, 0))
    sc.pack()
    assert len(sc.textures) == 2 and sc.globals["SkyboxTextureIndex"] != 0xFFFFFFFF
    with pytest.raises(OSError):
        sc.load_texture(str(tmp_path / "m.obj"), S.TEXTURE_RAW)


@pytest.mark.parametrize("mode", ["L", "LA", "RGB", "RGBA", "P"])
@pytest.mark.parametrize("rle", [False, True])
@pytest.mark.parametrize("top_down", [False, True])
def test_tga_matches_stb_image(tmp_path, mode, rle, top_down):
    w, h = 37, 23
    a = picture(w, h, 4)
    a[:, :9] = a[0, 0]                                        # runs for the run-length coder
    f = tmp_path / "t.tga"
    if mode == "P":
        im = PIL.fromarray(a[..., :3]).quantize(19)
    else:
        im = PIL.fromarray(a if mode == "RGBA" else a[..., :3] if mode == "RGB" else a[..., 0] if mode == "L" else a[..., [0, 3]], mode)
    im.save(f, compression="tga_rle" if rle else None, orientation=1 if top_down else -1)
    px = same(f)
    assert px.shape == (h, w, 4)


def test_tga_16_bit_colour_and_16_bit_indices(tmp_path):
    """5-5-5 true colour (type 2, 16 bits; and run-length type 10) and a colour-mapped file with 16-bit indices and a 15-bit
    palette: written by hand."""
    rng = np.random.default_rng(5)
    w, h = 11, 7
    px16 = rng.integers(0, 1 << 15, size=(h, w)).astype("<u2")
    hdr = struct.pack("<BBBHHBHHHHBB", 0, 0, 2, 0, 0, 0, 0, 0, w, h, 16, 0)
    (tmp_path / "a.tga").write_bytes(hdr + px16.tobytes())
    same(tmp_path / "a.tga")
    body = b""
    for row in px16:                                          # one raw packet per row (w <= 128)
        body += bytes([w - 1]) + row.tobytes()
    hdr = struct.pack("<BBBHHBHHHHBB", 3, 0, 10, 0, 0, 0, 0, 0, w, h, 15, 0x20)
    (tmp_path / "b.tga").write_bytes(hdr + b"abc" + body)
    same(tmp_path / "b.tga")
    pal = rng.integers(0, 1 << 15, size=300).astype("<u2")
    idx = rng.integers(0, 300, size=(h, w)).astype("<u2")
    hdr = struct.pack("<BBBHHBHHHHBB", 0, 1, 1, 0, 300, 15, 0, 0, w, h, 16, 0)
    (tmp_path / "c.tga").write_bytes(hdr + pal.tobytes() + idx.tobytes())
    same(tmp_path / "c.tga")


@pytest.mark.parametrize("mode", ["RGB", "RGBA", "P", "L", "1"])
@pytest.mark.parametrize("size", [(37, 23), (5, 3), (64, 1)])
def test_bmp_matches_stb_image(tmp_path, mode, size):
    w, h = size
    a = picture(w, h, 6)
    f = tmp_path / "t.bmp"
    if mode == "P":
        PIL.fromarray(a[..., :3]).quantize(23).save(f)
    elif mode == "1":
        PIL.fromarray(a[..., 0] > 128).save(f)
    else:
        PIL.fromarray(a if mode == "RGBA" else a[..., :3] if mode == "RGB" else a[..., 0], mode).save(f)
    px = same(f)
    assert px.shape == (h, w, 4)


def test_bmp_16_bit_bitfields_4_bit_and_top_down(tmp_path):
    """Forms Pillow does not write: 16-bit 5-5-5 (BI_RGB) and 5-6-5 (BITFIELDS), 32-bit BI_RGB whose alpha bytes are all zero
    (means opaque) or not, 4-bit palettes, negative height (rows top-down), an OS/2 12-byte header."""
    rng = np.random.default_rng(8)
    w, h = 13, 6

    def bmp(bpp, pixels_rows, compress=0, masks=b"", palette=b"", height=None, hsz=40):
        rows = b"".join(r + b"\0" * (-len(r) % 4) for r in pixels_rows)
        off = 14 + hsz + len(masks) + len(palette)
        if hsz == 12:
            info = struct.pack("<IHHHH", 12, w, h if height is None else height, 1, bpp)
        else:
            info = struct.pack("<IiiHHIIiiII", hsz, w, h if height is None else height, 1, bpp, compress, len(rows), 2835, 2835, 0, 0)
            info += b"\0" * (hsz - len(info))
        return b"BM" + struct.pack("<IHHI", off + len(rows), 0, 0, off) + info + masks + palette + rows
    v16 = rng.integers(0, 1 << 16, size=(h, w)).astype("<u2")
    (tmp_path / "a.bmp").write_bytes(bmp(16, [r.tobytes() for r in v16]))
    same(tmp_path / "a.bmp")
    (tmp_path / "b.bmp").write_bytes(bmp(16, [r.tobytes() for r in v16], compress=3, masks=struct.pack("<III", 0xF800, 0x07E0, 0x001F)))
    same(tmp_path / "b.bmp")
    v32 = rng.integers(0, 256, size=(h, w, 4)).astype(np.uint8)
    (tmp_path / "c.bmp").write_bytes(bmp(32, [r.tobytes() for r in v32]))
    same(tmp_path / "c.bmp")
    v32[..., 3] = 0
    (tmp_path / "d.bmp").write_bytes(bmp(32, [r.tobytes() for r in v32], height=-h))
    px = same(tmp_path / "d.bmp")
    assert (px[..., 3] == 1.0).all()
    pal = rng.integers(0, 256, size=(16, 4)).astype(np.uint8)
    idx = rng.integers(0, 16, size=(h, w + (w & 1)))
    rows4 = [bytes(int(r[i]) << 4 | int(r[i + 1]) for i in range(0, len(r), 2)) for r in idx]
    (tmp_path / "e.bmp").write_bytes(bmp(4, rows4, palette=pal.tobytes()))
    same(tmp_path / "e.bmp")
    v24 = rng.integers(0, 256, size=(h, w, 3)).astype(np.uint8)
    (tmp_path / "f.bmp").write_bytes(bmp(24, [r.tobytes() for r in v24], hsz=12))
    same(tmp_path / "f.bmp")
    pal3 = rng.integers(0, 256, size=(256, 3)).astype(np.uint8)
    idx8 = rng.integers(0, 252, size=(h, w)).astype(np.uint8)      # (stb_image reads only 252 entries of a 12-byte-header file's palette)
    (tmp_path / "g.bmp").write_bytes(bmp(8, [r.tobytes() for r in idx8], palette=pal3.tobytes(), hsz=12))
    same(tmp_path / "g.bmp")


def patch_adobe_transform(data, value):
    """Sets (or inserts, after SOI) an Adobe APP14 marker with the given colour-transform byte."""
    at = data.find(b"\xff\xee")
    if at >= 0 and data[at + 4:at + 9] == b"Adobe":
        return data[:at + 15] + bytes([value]) + data[at + 16:]
    return data[:2] + b"\xff\xee\x00\x0eAdobe\x00\x64\x00\x00\x00\x00" + bytes([value]) + data[2:]


def strip_jfif(data):
    at = data.find(b"\xff\xe0")
    if at < 0 or data[at + 4:at + 8] != b"JFIF":
        return data
    n = struct.unpack(">H", data[at + 2:at + 4])[0]
    return data[:at] + data[at + 2 + n:]


@pytest.mark.parametrize("case", ["cmyk", "ycck", "four-unknown-transform", "rgb-by-adobe-no-jfif", "adobe-0-with-jfif", "adobe-1", "ids-RGB"])
def test_jpeg_adobe_marker_and_four_components(tmp_path, case):
    """stb_image: 4 components are CMYK (Adobe transform 0), YCCK (2) or YCbCr + an ignored channel; 3 components are taken as RGB
    when their ids spell RGB or when an Adobe marker says 0 and there is no JFIF marker (stb_image.h:3878, 3955-3975)."""
    a = picture(45, 31, 3)
    f = tmp_path / "t.jpg"
    if case in ("cmyk", "ycck", "four-unknown-transform"):
        PIL.fromarray(a, "CMYK").save(f, quality=85)
        data = f.read_bytes()
        data = patch_adobe_transform(data, {"cmyk": 0, "ycck": 2, "four-unknown-transform": 1}[case])
    else:
        PIL.fromarray(a[..., :3]).save(f, quality=85, subsampling=0)
        data = f.read_bytes()
        if case == "rgb-by-adobe-no-jfif":
            data = patch_adobe_transform(strip_jfif(data), 0)
        elif case == "adobe-0-with-jfif":
            data = patch_adobe_transform(data, 0)
            assert b"JFIF" in data
        elif case == "adobe-1":
            data = patch_adobe_transform(strip_jfif(data), 1)
        else:                                                  # component ids 1, 2, 3 -> 'R', 'G', 'B' in SOF0 and SOS
            sof = data.find(b"\xff\xc0")
            sos = data.find(b"\xff\xda")
            d = bytearray(data)
            for k, c in enumerate(b"RGB"):
                assert d[sof + 10 + 3 * k] == k + 1 and d[sos + 5 + 2 * k] == k + 1
                d[sof + 10 + 3 * k] = c
                d[sos + 5 + 2 * k] = c
            data = bytes(d)
    f.write_bytes(data)
    same(f)


def test_random_valid_files_decode_like_stb_image():
    """A slice of tools/fuzz_image_files.py: random pictures written with random encoder settings (PNG, JPEG incl. CMYK, TGA, BMP)."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import fuzz_image_files
    assert fuzz_image_files.main(["--files", "300", "--seed", "5"]) == 0


@pytest.mark.parametrize("case", ["ppm8", "pgm8", "ppm16", "pgm16", "maxval-100", "odd-blanks", "comments", "maxval-256"])
def test_pnm_matches_stb_image(tmp_path, case):
    """stb_image's P5 / P6 reader (stb_image.h:7491-7618): the maximum value chooses the sample size and nothing else, two-byte
    samples are taken in host byte order and reduced to their high byte."""
    rng = np.random.default_rng(3)
    px = rng.integers(0, 256, size=(5, 7, 3), dtype=np.uint8)
    px16 = rng.integers(0, 65536, size=(5, 7, 3)).astype(">u2")
    data = {"ppm8": b"P6\n7 5\n255\n" + px.tobytes(),
            "pgm8": b"P5\n7 5\n255\n" + px[..., 0].tobytes(),
            "ppm16": b"P6\n7 5\n65535\n" + px16.tobytes(),
            "pgm16": b"P5\n7 5\n65535\n" + px16[..., 0].tobytes(),
            "maxval-100": b"P6 7 5 100\n" + (px // 3).tobytes(),
            "odd-blanks": b"P6\n7\t5\v\f255 " + px.tobytes(),
            "comments": b"P5#a\n# b\r7 5 # c\n255\n" + px[..., 0].tobytes(),
            "maxval-256": b"P5\n7 5\n256\n" + px16[..., 0].tobytes()}[case]
    f = tmp_path / "t.pnm"
    f.write_bytes(data)
    out = same(f)
    if case == "ppm8":
        assert np.allclose(out[0, 0, :3], (px[0, 0] / 255.0) ** 2.2, atol=1e-6) and out[0, 0, 3] == 1.0


@pytest.mark.parametrize("data", [b"P6\n7 5\n255\n" + bytes(104), b"P6\n0 5\n255\n" + bytes(105), b"P6\n7 5\n70000\n" + bytes(210), b"P6\n99999999999 5\n255\n"])
def test_pnm_refused_like_stb_image(tmp_path, data):
    f = tmp_path / "t.pnm"
    f.write_bytes(data)
    assert H.stbi_loadf(str(f)) is None and S.decode_image(str(f)) is None


def gif_parts(data):
    """(logical screen descriptor flags offset, global table slice, offset of the first image descriptor) of a GIF."""
    flags = data[10]
    table = slice(13, 13 + (3 * (2 << (flags & 7)) if flags & 0x80 else 0))
    at = table.stop
    while data[at] == 0x21:                                     # extensions: label, then sub-blocks
        at += 2
        while data[at]:
            at += 1 + data[at]
        at += 1
    assert data[at] == 0x2C
    return table, at


@pytest.mark.parametrize("case", ["palette-17", "palette-256", "grey", "two-colours", "transparent", "interlaced", "interlaced-odd", "animated",
                                  "bigger-screen-background", "bigger-screen-no-background", "local-palette", "local-palette-transparent",
                                  "truncated", "ends-before-image", "gif87a"])
def test_gif_first_frame_matches_stb_image(tmp_path, case):
    """stbi__gif_load: the first frame composed as stb_image composes it - untouched pixels (0, 0, 0, 0) or the background entry
    copied blue-first, transparent pixels not drawn, local palettes, interlace."""
    a = picture(45, 30 if case != "interlaced-odd" else 29, 3)
    f = tmp_path / "t.gif"
    rgb = PIL.fromarray(a[..., :3])
    if case == "palette-256":
        rgb.quantize(256).save(f)
    elif case == "grey":
        PIL.fromarray(a[..., 0]).save(f)
    elif case == "two-colours":
        rgb.quantize(2).save(f)
    elif case in ("transparent", "local-palette-transparent"):
        rgb.quantize(5).save(f, transparency=2)
    elif case in ("interlaced", "interlaced-odd"):
        rgb.quantize(64).save(f, interlace=True)
        assert f.read_bytes()[gif_parts(f.read_bytes())[1] + 9] & 0x40
    elif case == "animated":
        frames = [PIL.fromarray(picture(45, 30, k)[..., :3]).quantize(30) for k in range(3)]
        frames[0].save(f, save_all=True, append_images=frames[1:], duration=40, loop=0)
    else:
        rgb.quantize(17).save(f)
    data = bytearray(f.read_bytes())
    if case.startswith("bigger-screen"):
        data[6:10] = struct.pack("<HH", 45 + 6, 30 + 4)         # the frame keeps its 45 x 30 rectangle at (0, 0)
        data[11] = 3 if case == "bigger-screen-background" else 0
    elif case.startswith("local-palette"):
        table, at = gif_parts(bytes(data))
        pal = bytes(data[table])
        size_bits = data[10] & 7
        desc = bytearray(data[at:at + 10])
        desc[9] |= 0x80 | size_bits
        data = data[:10] + bytes([data[10] & 0x7F]) + data[11:13] + data[table.stop:at] + desc + pal + data[at + 10:]
    elif case == "truncated":
        data = data[:len(data) // 2]
    elif case == "ends-before-image":
        table, at = gif_parts(bytes(data))
        data = data[:at] + b";"
    elif case == "gif87a":
        data[3:6] = b"87a"
    f.write_bytes(bytes(data))
    if case == "ends-before-image":
        assert H.stbi_loadf(str(f)) is None and S.decode_image(str(f)) is None
        return
    out = same(f)
    if case == "bigger-screen-no-background":
        assert (out[:, 45:] == 0).all() and (out[30:] == 0).all()
    if case == "bigger-screen-background":
        assert (out[:, 45:, 3] == 1).all()


def packbits(row):
    """PackBits with literal runs only where nothing repeats (enough to exercise both packet kinds)."""
    out, i = bytearray(), 0
    while i < len(row):
        j = i
        while j + 1 < len(row) and row[j + 1] == row[i] and j - i < 127:
            j += 1
        if j > i:
            out += bytes([257 - (j - i + 1), row[i]])
            i = j + 1
        else:
            k = i
            while k + 1 < len(row) and row[k + 1] != row[k] and k - i < 127:
                k += 1
            out += bytes([k - i]) + bytes(row[i:k + 1])
            i = k + 1
    return bytes(out)


def write_psd(path, planes, depth=8, rle=False, mode=3, noop=False):
    """planes: (channels, h, w) array of uint8 (depth 8) or uint16 (depth 16)."""
    c, h, w = planes.shape
    head = b"8BPS" + struct.pack(">H6xHIIHH", 1, c, h, w, depth, mode) + struct.pack(">I", 0) + struct.pack(">I", 4) + b"abcd" + struct.pack(">I", 0)
    if rle:
        rows = [packbits(planes[k, y].astype(np.uint8).tobytes()) for k in range(c) for y in range(h)]
        if noop:
            rows[0] = b"\x80" + rows[0]
        body = struct.pack(">H", 1) + b"".join(struct.pack(">H", len(r)) for r in rows) + b"".join(rows)
    else:
        body = struct.pack(">H", 0) + (planes.astype(">u2") if depth == 16 else planes.astype(np.uint8)).tobytes()
    with open(path, "wb") as f:
        f.write(head + body)


@pytest.mark.parametrize("channels", [1, 2, 3, 4, 5])
@pytest.mark.parametrize("kind", ["raw8", "raw16", "rle8", "rle8-noop"])
def test_psd_matches_stb_image(tmp_path, channels, kind):
    """stbi__psd_load: planar channels, the first four used, missing ones filled, 16-bit samples reduced to their high byte, PackBits
    rows, and the white un-matting of colour under partial alpha."""
    rng = np.random.default_rng(channels)
    h, w = 13, 21
    if kind == "raw16":
        planes = rng.integers(0, 65536, size=(channels, h, w)).astype(np.uint16)
    else:
        planes = rng.integers(0, 256, size=(channels, h, w)).astype(np.uint8)
        planes[:, 3:6, :] = planes[:, 3:6, :1]                  # flat rows: replicate packets
        if channels >= 4:
            planes[3, 0, :4] = (0, 255, 1, 254)                 # alpha extremes next to the un-matting cases
    f = tmp_path / "t.psd"
    write_psd(f, planes, depth=16 if kind == "raw16" else 8, rle=kind.startswith("rle"), noop=kind.endswith("noop"))
    same(f)


@pytest.mark.parametrize("what", ["cmyk-mode", "depth-32", "version-2", "seventeen-channels", "bad-compression", "rle-overrun"])
def test_psd_refused_like_stb_image(tmp_path, what):
    planes = np.zeros((3, 4, 5), np.uint8)
    f = tmp_path / "t.psd"
    write_psd(f, planes, mode=4 if what == "cmyk-mode" else 3)
    d = bytearray(f.read_bytes())
    if what == "depth-32":
        d[22:24] = struct.pack(">H", 32)
    elif what == "version-2":
        d[4:6] = struct.pack(">H", 2)
    elif what == "seventeen-channels":
        d[12:14] = struct.pack(">H", 17)
    elif what == "bad-compression":
        d[42:44] = struct.pack(">H", 2)
    elif what == "rle-overrun":
        write_psd(f, planes, rle=True)
        d = bytearray(f.read_bytes())
        d[44 + 2 * 12] = 100                                   # a literal packet longer than the channel
    f.write_bytes(bytes(d))
    assert H.stbi_loadf(str(f)) is None and S.decode_image(str(f)) is None


def write_pic(path, rgba, packets, rng):
    """packets: list of (type, channel mask).  Rows are written packet by packet; type 1 / 2 with runs found in the data."""
    h, w = rgba.shape[:2]
    head = bytearray(104)
    head[0:4] = b"\x53\x80\xf6\x34"
    head[88:92] = b"PICT"
    head[92:96] = struct.pack(">HH", w, h)
    head[96:100] = struct.pack(">f", 1.0)
    head[100:102] = struct.pack(">H", 3)
    body = bytearray()
    for k, (ptype, mask) in enumerate(packets):
        body += bytes([1 if k + 1 < len(packets) else 0, 8, ptype, mask])
    chans = lambda mask: [i for i in range(4) if mask & (0x80 >> i)]
    for y in range(h):
        for ptype, mask in packets:
            px = [bytes(rgba[y, x, chans(mask)].tolist()) for x in range(w)]
            x = 0
            while x < w:
                run = 1
                while x + run < w and px[x + run] == px[x] and run < 200:
                    run += 1
                if ptype == 0:
                    body += px[x]
                    x += 1
                elif ptype == 1:
                    body += bytes([run]) + px[x]
                    x += run
                else:
                    if run >= 3:
                        if run > 128 or rng.random() < 0.2:
                            body += bytes([128]) + struct.pack(">H", run) + px[x]
                        else:
                            body += bytes([run + 127]) + px[x]
                        x += run
                    else:
                        n = 1
                        while x + n < w and n < 100 and not (x + n + 2 < w and px[x + n] == px[x + n + 1] == px[x + n + 2]):
                            n += 1
                        body += bytes([n - 1]) + b"".join(px[x:x + n])
                        x += n
    with open(path, "wb") as f:
        f.write(bytes(head) + bytes(body))


@pytest.mark.parametrize("packets", [[(0, 0xE0)], [(1, 0xE0)], [(2, 0xE0)], [(2, 0xE0), (2, 0x10)], [(0, 0xF0)], [(1, 0x80), (2, 0x40), (0, 0x20)],
                                     [(2, 0x10)], [(2, 0xC0), (1, 0x30)]])
def test_pic_matches_stb_image(tmp_path, packets):
    """stbi__pic_load: raw, pure run-length and mixed run-length packets, channels spread over packets, missing channels 255."""
    rng = np.random.default_rng(len(packets))
    a = picture(150, 9, 5)
    a[2:5] = a[2:5, :1]                                        # flat rows: long runs (16-bit counts in mixed packets)
    a[6, 10:40] = a[6, 10]
    f = tmp_path / "t.pic"
    write_pic(f, a, packets, rng)
    out = same(f)
    mask = 0
    for _, m in packets:
        mask |= m
    for i in range(4):
        if not mask & (0x80 >> i):
            assert (out[..., i] == 1.0).all()
