#!/usr/bin/env python
"""Differential fuzz of the image decoders (host/image.cpp) against the reference's vendored stb_image on VALID random files.  CPU only.

tests/test_image_decode.py covers the formats case by case; tools/fuzz_image_decoders.py mutates files for memory safety.  This
tool writes well-formed files with random content AND random encoder settings through Pillow - PNG (all modes, bit depths 1-16,
palettes of odd sizes, tRNS, interlace is not writable by Pillow and stays with the unit tests, compression levels, odd sizes),
JPEG (quality 1-100, 4:4:4 / 4:2:2 / 4:2:0 / 4:1:1 / 4:4:0 subsampling, progressive, optimised tables, greyscale, CMYK and YCCK
with the Adobe marker, restart intervals, sizes around the MCU edges), TGA (RLE or not, palettes, 8/24/32 bit), BMP (1/4/8/24/32
bit), binary PGM / PPM (8 and 16 bit, hand-written headers), GIF (palette sizes, transparency, interlace, animation, larger
screens with and without a background index, local palettes), PSD (1-6 channels, 8 / 16 bit, raw / PackBits), Softimage PIC (raw / run-length / mixed packets) and noisy as well as smooth pictures - and demands that `stbi_loadf(path, ..., 4)` of the reference (oracle/_ref/
libref_host.so) and DecodeImageFile return the same float pixels bit for bit.

    python tools/fuzz_image_files.py [--files 1500] [--seed 1]

Exit code 1 and the path of the file on the first difference."""
import argparse
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def random_picture(rng, w, h):
    kind = rng.integers(0, 4)
    y, x = np.mgrid[0:h, 0:w]
    if kind == 0:                                              # noise: every coefficient of every block is busy
        return rng.integers(0, 256, size=(h, w, 4)).astype(np.uint8)
    if kind == 1:                                              # smooth gradients: long zero runs, end-of-block everywhere
        f = rng.uniform(0.5, 6.0, size=4)
        return np.stack([127.5 + 127.5 * np.sin(x / f[0] + y / f[1]), 127.5 + 127.5 * np.cos(y / f[2]), (x * 255 // max(1, w - 1)) + 0 * y,
                         127.5 + 127.5 * np.sin((x + y) / f[3])], -1).astype(np.uint8)
    if kind == 2:                                              # flat areas with hard edges
        q = int(rng.integers(2, 17))
        return (rng.integers(0, 256, size=(h // q + 1, w // q + 1, 4)).astype(np.uint8).repeat(q, 0).repeat(q, 1))[:h, :w]
    a = rng.integers(0, 256, size=(h, w, 4)).astype(np.uint8)  # extremes: saturated values next to each other
    return np.where(a > 127, 255, 0).astype(np.uint8)


def write_random_file(rng, d, k):
    from PIL import Image
    w = int(rng.choice([1, 2, 7, 8, 9, 15, 16, 17, 31, 33, 64, 100])) if rng.random() < 0.7 else int(rng.integers(1, 200))
    h = int(rng.choice([1, 2, 7, 8, 9, 15, 16, 17, 31, 33, 64, 100])) if rng.random() < 0.7 else int(rng.integers(1, 200))
    a = random_picture(rng, w, h)
    fmt = rng.choice(["png", "jpg", "jpg", "tga", "bmp", "pnm", "gif", "psd", "pic"])
    path = os.path.join(d, "f%05d.%s" % (k, fmt))
    if fmt == "png":
        mode = rng.choice(["L", "LA", "RGB", "RGBA", "P", "P-trns", "1", "I;16", "RGB-trns", "L-trns"])
        kw = dict(compress_level=int(rng.integers(0, 10)), optimize=bool(rng.integers(0, 2)))
        if mode == "P":
            Image.fromarray(a[..., :3]).quantize(int(rng.integers(2, 257))).save(path, **kw)
        elif mode == "P-trns":
            n = int(rng.integers(2, 200))
            Image.fromarray(a[..., :3]).quantize(n).save(path, transparency=bytes(rng.integers(0, 256, size=int(rng.integers(1, n + 1))).tolist()), **kw)
        elif mode == "1":
            Image.fromarray(a[..., 0] > 128).save(path, **kw)
        elif mode == "I;16":
            Image.fromarray((a[..., 0].astype(np.uint16) * 257 + a[..., 1]) & 0xFFFF).save(path, **kw)
        elif mode == "RGB-trns":
            Image.fromarray(a[..., :3]).save(path, transparency=tuple(int(v) for v in a[0, 0, :3]), **kw)
        elif mode == "L-trns":
            Image.fromarray(a[..., 0]).save(path, transparency=int(a[0, 0, 0]), **kw)
        else:
            Image.fromarray(a if mode == "RGBA" else a[..., :3] if mode == "RGB" else a[..., 0] if mode == "L" else a[..., [0, 3]], mode).save(path, **kw)
    elif fmt == "jpg":
        mode = rng.choice(["RGB", "RGB", "L", "CMYK"])
        im = Image.fromarray(a[..., :3]) if mode == "RGB" else Image.fromarray(a[..., 0]) if mode == "L" else Image.fromarray(a, "CMYK")
        kw = dict(quality=int(rng.choice([1, 5, 25, 50, 75, 90, 95, 100])), progressive=bool(rng.integers(0, 2)), optimize=bool(rng.integers(0, 2)))
        if mode == "RGB":
            kw["subsampling"] = rng.choice(["4:4:4", "4:2:2", "4:2:0", "4:1:1", "4:4:0"]) if rng.random() < 0.8 else int(rng.integers(0, 3))
        if rng.random() < 0.4:
            kw["restart_marker_blocks"] = int(rng.integers(1, 9))
        try:
            im.save(path, **kw)
        except Exception:                                      # a Pillow without some option: plain settings
            im.save(path, quality=kw["quality"])
    elif fmt == "tga":
        mode = rng.choice(["L", "LA", "RGB", "RGBA", "P"])
        im = (Image.fromarray(a[..., :3]).quantize(int(rng.integers(2, 257))) if mode == "P" else
              Image.fromarray(a if mode == "RGBA" else a[..., :3] if mode == "RGB" else a[..., 0] if mode == "L" else a[..., [0, 3]], mode))
        kw = {}
        if rng.random() < 0.5:
            kw["compression"] = "tga_rle"
        if rng.random() < 0.5:
            kw["orientation"] = int(rng.choice([1, -1]))
        im.save(path, **kw)
    elif fmt == "gif":                                         # first frame; sometimes a larger screen, a background index, a local palette
        import struct
        kw = {}
        if rng.random() < 0.4:
            kw["interlace"] = True
        n = int(rng.integers(2, 257))
        if rng.random() < 0.4:
            kw["transparency"] = int(rng.integers(0, n))
        im = Image.fromarray(a[..., 0]) if rng.random() < 0.2 else Image.fromarray(a[..., :3]).quantize(n)
        if rng.random() < 0.2:
            im.save(path, save_all=True, append_images=[Image.fromarray(random_picture(rng, w, h)[..., :3]).quantize(n)], duration=30, **kw)
        else:
            im.save(path, **kw)
        data = bytearray(open(path, "rb").read())
        if rng.random() < 0.4:
            data[6:10] = struct.pack("<HH", w + int(rng.integers(0, 9)), h + int(rng.integers(0, 9)))
            data[11] = int(rng.integers(0, 2 << (data[10] & 7))) if (data[10] & 0x80) and rng.random() < 0.6 else 0
        if rng.random() < 0.3 and (data[10] & 0x80):            # move the global palette behind the image descriptor
            table = slice(13, 13 + 3 * (2 << (data[10] & 7)))
            at = table.stop
            while data[at] == 0x21:
                at += 2
                while data[at]:
                    at += 1 + data[at]
                at += 1
            if data[at] == 0x2C:
                desc = bytearray(data[at:at + 10])
                desc[9] |= 0x80 | (data[10] & 7)
                data = data[:10] + bytes([data[10] & 0x7F]) + data[11:13] + data[table.stop:at] + desc + bytes(data[table]) + data[at + 10:]
        open(path, "wb").write(bytes(data))
    elif fmt == "psd":                                         # flattened RGB image written by hand (tests/test_image_decode.py write_psd)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import test_image_decode as T
        c = int(rng.integers(1, 7))
        deep = rng.random() < 0.3
        planes = (rng.integers(0, 65536, size=(c, h, w)).astype(np.uint16) if deep else
                  np.stack([random_picture(rng, w, h)[..., k % 4] for k in range(c)]).astype(np.uint8))
        T.write_psd(path, planes, depth=16 if deep else 8, rle=(not deep) and rng.random() < 0.6, noop=rng.random() < 0.2)
    elif fmt == "pic":                                         # Softimage PIC written by hand (tests/test_image_decode.py write_pic)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import test_image_decode as T
        masks = [[0xE0], [0xF0], [0xE0, 0x10], [0x80, 0x40, 0x20], [0xC0, 0x30], [0x10], [0x80, 0x20]][rng.integers(0, 7)]
        T.write_pic(path, a, [(int(rng.integers(0, 3)), m) for m in masks], rng)
    elif fmt == "pnm":                                         # binary P5 / P6, 8 and 16 bit, header written by hand
        grey, wide = bool(rng.integers(0, 2)), bool(rng.integers(0, 2))
        body = (rng.integers(0, 65536, size=(h, w, 1 if grey else 3)).astype(">u2") if wide else a[..., :1] if grey else a[..., :3]).tobytes()
        sep = [b"\n", b" ", b"\t", b"\r\n", b"\n# c\n"]
        hdr = (b"P5" if grey else b"P6") + sep[rng.integers(0, 5)] + str(w).encode() + sep[rng.integers(0, 5)] + str(h).encode() + sep[rng.integers(0, 5)]
        hdr += str(int(rng.integers(256, 65536)) if wide else int(rng.integers(1, 256))).encode() + sep[rng.integers(0, 3)][:1]
        open(path, "wb").write(hdr + body)
    else:
        mode = rng.choice(["RGB", "RGBA", "L", "P", "1"])
        im = (Image.fromarray(a[..., :3]).quantize(int(rng.integers(2, 257))) if mode == "P" else Image.fromarray(a[..., 0] > 128) if mode == "1" else
              Image.fromarray(a if mode == "RGBA" else a[..., :3] if mode == "RGB" else a[..., 0], mode))
        im.save(path)
    return path


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--files", type=int, default=1500)
    ap.add_argument("--seed", type=int, default=1)
    args = ap.parse_args(argv)
    from oracle.ref import host_binding as H
    from path_tracer_b200 import scene as S
    assert H.available(), "oracle/_ref/libref_host.so cannot be built here"
    H.lib()
    d = tempfile.mkdtemp(prefix="fuzz_img_")
    rng = np.random.default_rng(args.seed)
    counts, unread = {}, 0
    for k in range(args.files):
        path = write_random_file(rng, d, k)
        ref, ours = H.stbi_loadf(path), S.decode_image(path)
        ext = path.rsplit(".", 1)[1]
        if ref is None or ours is None:
            if (ref is None) != (ours is None):
                print(path, ": read by one decoder only (ours: %s, stb_image: %s)" % (ours is not None, ref is not None))
                return 1
            unread += 1
            os.remove(path)
            continue
        if ours.shape != ref.shape or ours.tobytes() != ref.tobytes():
            if ours.shape == ref.shape:
                bad = np.argwhere((ours.view(np.uint32) != ref.view(np.uint32)).any(axis=-1))
                print(path, ": %d of %d pixels differ, first at %s: ours %s, stb_image %s" % (len(bad), ours.shape[0] * ours.shape[1], bad[0], ours[tuple(bad[0])], ref[tuple(bad[0])]))
            else:
                print(path, ": shapes differ", ours.shape, ref.shape)
            return 1
        counts[ext] = counts.get(ext, 0) + 1
        os.remove(path)
        if (k + 1) % 250 == 0:
            print("%d files: equal %s, %d read by neither" % (k + 1, counts, unread), flush=True)
    print("done: %d files, bit-identical with stb_image: %s; read by neither: %d" % (args.files, counts, unread))
    return 0


if __name__ == "__main__":
    sys.exit(main())
