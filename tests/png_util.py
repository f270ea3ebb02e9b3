"""Minimal PNG writer (8-bit RGB / RGBA, no filtering) for synthetic anno datasets: the reference CLIs read their images
and *_mask.png label files through dlib's bundled libpng (annonet.cpp:146-160)."""
import struct
import zlib

import numpy as np


def write_png(path, img):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w, c = img.shape
    assert c in (3, 4)
    raw = b"".join(b"\x00" + img[y].tobytes() for y in range(h))

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n")
        f.write(chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2 if c == 3 else 6, 0, 0, 0)))
        f.write(chunk(b"IDAT", zlib.compress(raw, 6)))
        f.write(chunk(b"IEND", b""))


# the reference's default anno classes (annonet_parse_anno_classes.cpp:22-28); rgba (0, 0, 0, 0) = ignore
DEFAULT_CLASS_COLORS = [(0, 255, 0, 64), (255, 255, 0, 128), (255, 0, 0, 128)]


def write_anno_dataset(directory, rng, count, h, w, synth_image):
    """count images <i>.png with masks <i>.png_mask.png: blocky class regions, a band of ignored pixels"""
    import os
    for i in range(count):
        img = synth_image(rng, h, w)
        yy, xx = np.mgrid[0:h, 0:w]
        lab = ((yy // 40) + (xx // 40) + i) % 3
        mask = np.zeros((h, w, 4), np.uint8)
        for k, col in enumerate(DEFAULT_CLASS_COLORS):
            mask[lab == k] = col
            img[lab == k, k] = np.clip(img[lab == k, k].astype(np.int32) + 40, 0, 255)   # make the classes learnable
        mask[:, w // 2 - 3:w // 2 + 3] = (0, 0, 0, 0)
        write_png(os.path.join(directory, f"{i}.png"), img)
        write_png(os.path.join(directory, f"{i}.png_mask.png"), mask)


def read_png(path):
    """8-bit RGB / RGBA PNG -> array (handles the five scanline filters; enough for dlib's save_png output)"""
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h, ctype = 8, b"", 0, 0, 0
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        if tag == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
            assert depth == 8 and ctype in (2, 6)
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    c = 3 if ctype == 2 else 4
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + w * c)
    out = np.zeros((h, w * c), np.int32)
    for y in range(h):
        f, line = int(raw[y, 0]), raw[y, 1:].astype(np.int32)
        prev = out[y - 1] if y else np.zeros(w * c, np.int32)
        if f == 0:
            out[y] = line
        elif f == 2:
            out[y] = (line + prev) & 255
        else:
            cur = np.zeros(w * c, np.int32)
            for x in range(w * c):
                a = cur[x - c] if x >= c else 0
                b = prev[x]
                cc = prev[x - c] if x >= c else 0
                if f == 1:
                    p = a
                elif f == 3:
                    p = (a + b) // 2
                else:
                    pa, pb, pc = abs(b - cc), abs(a - cc), abs(a + b - 2 * cc)
                    p = a if (pa <= pb and pa <= pc) else (b if pb <= pc else cc)
                cur[x] = (line[x] + p) & 255
            out[y] = cur
    return out.reshape(h, w, c).astype(np.uint8)
