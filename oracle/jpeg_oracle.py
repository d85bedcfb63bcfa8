"""CPU oracle for the JPEG re-encode of the blurred image -- TEST INFRASTRUCTURE, NOT PRODUCT
(same import rule as the other oracle modules).

PARITY PINNED.  The reference serves ``GET /fetch/contents`` by saving the blurred PIL image as
JPEG with Pillow's defaults (main.py:100-103 ``image.save(img_io, 'JPEG')``; the same call in
src/utils.py:12-16 ``encode_image``): baseline sequential JPEG, quality 75, 4:2:0 chroma
subsampling, the standard Huffman tables of ITU-T T.81 Annex K, no restart markers, a JFIF 1.01
header.  The arithmetic lives in libjpeg(-turbo) behind Pillow, and Pillow IS installed here
(12.2.0, libjpeg-turbo), so this numpy restatement of libjpeg's encoder --

    jccolor.c   rgb_ycc_convert   16-bit fixed point colour conversion
    jcsample.c  h2v2_downsample   2x2 box filter, rounding bias alternating 1, 2
    jcprepct.c / jcsample.c       right / bottom edge replication
    jfdctint.c  jpeg_fdct_islow   13-bit LL&M integer forward DCT, output scaled by 8
    jcdctmgr.c  quantize          round-half-away division by 8 * q
    jccoefct.c  compress_data     dummy blocks (zero AC, DC of the previous block) beyond the edges
    jchuff.c    encode_one_block  DC differences, run/size symbols, ZRL, EOB, 0xFF stuffing

-- is checked byte for byte against the real ``Image.save(..., 'JPEG')`` on every run of
tests/test_jpeg_oracle.py (random, smooth and flat images; sizes that are and are not multiples
of 16; qualities 30..95), and the GPU tests compare the CUDA bytes with live Pillow directly.
"""
from __future__ import annotations

import numpy as np

# ITU-T T.81 Annex K.1 / K.2 base tables (natural order), as in libjpeg's jcparam.c
STD_LUMA_Q = np.array([
    16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55,
    14, 13, 16, 24, 40, 57, 69, 56, 14, 17, 22, 29, 51, 87, 80, 62,
    18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92,
    49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99], dtype=np.int64)
STD_CHROMA_Q = np.array([
    17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99,
    24, 26, 56, 99, 99, 99, 99, 99, 47, 66, 99, 99, 99, 99, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
    99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99], dtype=np.int64)

# jpeg_natural_order: zigzag position k -> natural (row-major) index
ZIGZAG = np.array([
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5,
    12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
    58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63], dtype=np.int64)

# Annex K.3 - K.6 Huffman tables: (bits[1..16], values)
DC_LUMA = ([0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0], list(range(12)))
DC_CHROMA = ([0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0], list(range(12)))
AC_LUMA = ([0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 125], [
    1, 2, 3, 0, 4, 17, 5, 18, 33, 49, 65, 6, 19, 81, 97, 7, 34, 113, 20, 50, 129, 145, 161, 8, 35, 66, 177, 193, 21, 82,
    209, 240, 36, 51, 98, 114, 130, 9, 10, 22, 23, 24, 25, 26, 37, 38, 39, 40, 41, 42, 52, 53, 54, 55, 56, 57, 58, 67,
    68, 69, 70, 71, 72, 73, 74, 83, 84, 85, 86, 87, 88, 89, 90, 99, 100, 101, 102, 103, 104, 105, 106, 115, 116, 117,
    118, 119, 120, 121, 122, 131, 132, 133, 134, 135, 136, 137, 138, 146, 147, 148, 149, 150, 151, 152, 153, 154, 162,
    163, 164, 165, 166, 167, 168, 169, 170, 178, 179, 180, 181, 182, 183, 184, 185, 186, 194, 195, 196, 197, 198, 199,
    200, 201, 202, 210, 211, 212, 213, 214, 215, 216, 217, 218, 225, 226, 227, 228, 229, 230, 231, 232, 233, 234, 241,
    242, 243, 244, 245, 246, 247, 248, 249, 250])
AC_CHROMA = ([0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 119], [
    0, 1, 2, 3, 17, 4, 5, 33, 49, 6, 18, 65, 81, 7, 97, 113, 19, 34, 50, 129, 8, 20, 66, 145, 161, 177, 193, 9, 35, 51,
    82, 240, 21, 98, 114, 209, 10, 22, 36, 52, 225, 37, 241, 23, 24, 25, 26, 38, 39, 40, 41, 42, 53, 54, 55, 56, 57, 58,
    67, 68, 69, 70, 71, 72, 73, 74, 83, 84, 85, 86, 87, 88, 89, 90, 99, 100, 101, 102, 103, 104, 105, 106, 115, 116,
    117, 118, 119, 120, 121, 122, 130, 131, 132, 133, 134, 135, 136, 137, 138, 146, 147, 148, 149, 150, 151, 152, 153,
    154, 162, 163, 164, 165, 166, 167, 168, 169, 170, 178, 179, 180, 181, 182, 183, 184, 185, 186, 194, 195, 196, 197,
    198, 199, 200, 201, 202, 210, 211, 212, 213, 214, 215, 216, 217, 218, 226, 227, 228, 229, 230, 231, 232, 233, 234,
    242, 243, 244, 245, 246, 247, 248, 249, 250])


def quant_tables(quality: int = 75):
    """jpeg_set_quality(quality, force_baseline=TRUE): (luma, chroma) in natural order."""
    quality = min(max(int(quality), 1), 100)
    scale = 5000 // quality if quality < 50 else 200 - quality * 2
    out = []
    for base in (STD_LUMA_Q, STD_CHROMA_Q):
        out.append(np.clip((base * scale + 50) // 100, 1, 255))
    return out


def derive_codes(bits, vals):
    """jpeg_make_c_derived_tbl: symbol -> (code, length)."""
    code, k, table = 0, 0, {}
    for length in range(1, 17):
        for _ in range(bits[length - 1]):
            table[vals[k]] = (code, length)
            code += 1
            k += 1
        code <<= 1
    return table


def rgb_to_ycc(rgb: np.ndarray):
    """jccolor.c rgb_ycc_convert on uint8 [H, W, 3] -> three int64 [H, W] planes in 0..255."""
    def fix(x):
        return int(x * 65536 + 0.5)
    r, g, b = (rgb[..., c].astype(np.int64) for c in range(3))
    half, off = 1 << 15, 128 << 16
    y = (fix(0.29900) * r + fix(0.58700) * g + fix(0.11400) * b + half) >> 16
    cb = (-fix(0.16874) * r - fix(0.33126) * g + fix(0.50000) * b + off + half - 1) >> 16
    cr = (fix(0.50000) * r - fix(0.41869) * g - fix(0.08131) * b + off + half - 1) >> 16
    return y, cb, cr


def _pad_right(plane: np.ndarray, cols: int) -> np.ndarray:
    if plane.shape[1] >= cols:
        return plane[:, :cols]
    return np.concatenate([plane, np.repeat(plane[:, -1:], cols - plane.shape[1], axis=1)], axis=1)


def _pad_bottom(plane: np.ndarray, rows: int) -> np.ndarray:
    if plane.shape[0] >= rows:
        return plane[:rows]
    return np.concatenate([plane, np.repeat(plane[-1:], rows - plane.shape[0], axis=0)], axis=0)


def component_planes(rgb: np.ndarray):
    """Sample planes as the DCT sees them (jcprepct.c + jcsample.c), 4:2:0.

    Returns (Y [16*mcu_rows, 8*y_blocks_w], Cb, Cr [8*mcu_rows, 8*c_blocks_w]) and the block
    geometry.  Full-resolution rows are first padded to an even count by repeating the last
    row, then down-sampled; only after that is every component padded to whole MCU rows by
    repeating ITS last row."""
    h, w = rgb.shape[:2]
    y, cb, cr = rgb_to_ycc(rgb)
    yb_w, cb_w = -(-w // 8), -(-w // 16)                      # width_in_blocks per component
    yb_h, cb_h = -(-h // 8), -(-h // 16)                      # height_in_blocks
    mcu_rows, mcu_cols = -(-h // 16), -(-w // 16)
    h_even = h + (h & 1)
    y = _pad_right(_pad_bottom(y, h_even), yb_w * 8)
    planes = [_pad_bottom(y, mcu_rows * 16)]
    bias = np.tile(np.array([1, 2], dtype=np.int64), cb_w * 4)[None, :]     # 1, 2, 1, 2, ... per output row
    for c in (cb, cr):
        c = _pad_right(_pad_bottom(c, h_even), cb_w * 16)
        s = c[0::2, 0::2] + c[0::2, 1::2] + c[1::2, 0::2] + c[1::2, 1::2]
        planes.append(_pad_bottom((s + bias) >> 2, mcu_rows * 8))
    geom = dict(yb_w=yb_w, yb_h=yb_h, cb_w=cb_w, cb_h=cb_h, mcu_rows=mcu_rows, mcu_cols=mcu_cols)
    return planes, geom


def fdct_islow(blocks: np.ndarray) -> np.ndarray:
    """jfdctint.c jpeg_fdct_islow on int64 [..., 8, 8] samples (already centred: sample - 128).
    Output coefficients are scaled up by 8, as libjpeg leaves them for the quantiser."""
    CONST_BITS, PASS1_BITS = 13, 2
    F = dict(c0_298=2446, c0_390=3196, c0_541=4433, c0_765=6270, c0_899=7373, c1_175=9633,
             c1_501=12299, c1_847=15137, c1_961=16069, c2_053=16819, c2_562=20995, c3_072=25172)

    def descale(x, n):
        return (x + (1 << (n - 1))) >> n

    def one_pass(d, first: bool):
        # d[..., k]: the 8 inputs of every 1-D transform along the last axis
        d0, d1, d2, d3, d4, d5, d6, d7 = (d[..., k] for k in range(8))
        tmp0, tmp7 = d0 + d7, d0 - d7
        tmp1, tmp6 = d1 + d6, d1 - d6
        tmp2, tmp5 = d2 + d5, d2 - d5
        tmp3, tmp4 = d3 + d4, d3 - d4
        tmp10, tmp13 = tmp0 + tmp3, tmp0 - tmp3
        tmp11, tmp12 = tmp1 + tmp2, tmp1 - tmp2
        if first:
            o0, o4 = (tmp10 + tmp11) << PASS1_BITS, (tmp10 - tmp11) << PASS1_BITS
            n = CONST_BITS - PASS1_BITS
        else:
            o0, o4 = descale(tmp10 + tmp11, PASS1_BITS), descale(tmp10 - tmp11, PASS1_BITS)
            n = CONST_BITS + PASS1_BITS
        z1 = (tmp12 + tmp13) * F["c0_541"]
        o2 = descale(z1 + tmp13 * F["c0_765"], n)
        o6 = descale(z1 + tmp12 * (-F["c1_847"]), n)
        z1, z2, z3, z4 = tmp4 + tmp7, tmp5 + tmp6, tmp4 + tmp6, tmp5 + tmp7
        z5 = (z3 + z4) * F["c1_175"]
        t4, t5, t6, t7 = tmp4 * F["c0_298"], tmp5 * F["c2_053"], tmp6 * F["c3_072"], tmp7 * F["c1_501"]
        z1, z2 = z1 * (-F["c0_899"]), z2 * (-F["c2_562"])
        z3, z4 = z3 * (-F["c1_961"]) + z5, z4 * (-F["c0_390"]) + z5
        o7, o5 = descale(t4 + z1 + z3, n), descale(t5 + z2 + z4, n)
        o3, o1 = descale(t6 + z2 + z3, n), descale(t7 + z1 + z4, n)
        return np.stack([o0, o1, o2, o3, o4, o5, o6, o7], axis=-1)

    rows = one_pass(blocks.astype(np.int64), True)                 # along each row
    cols = one_pass(np.swapaxes(rows, -1, -2), False)              # along each column
    return np.swapaxes(cols, -1, -2)


def quantize(coef: np.ndarray, qtbl: np.ndarray) -> np.ndarray:
    """jcdctmgr.c quantize: sign(c) * ((|c| + (8q >> 1)) // (8q)), q in natural order [64]."""
    q8 = (qtbl.reshape(8, 8) << 3).astype(np.int64)
    mag = (np.abs(coef) + (q8 >> 1)) // q8
    return np.where(coef < 0, -mag, mag)


def _blocks_of(plane: np.ndarray) -> np.ndarray:
    """[H8*8, W8*8] -> [H8, W8, 8, 8]."""
    h8, w8 = plane.shape[0] // 8, plane.shape[1] // 8
    return plane.reshape(h8, 8, w8, 8).swapaxes(1, 2)


def quantized_blocks(rgb: np.ndarray, quality: int = 75):
    """Quantised coefficients in ZIGZAG order, in the order the scan emits them:
    int64 [mcu_rows * mcu_cols, 6, 64] (Y0 Y1 Y2 Y3 Cb Cr per MCU), dummy blocks included."""
    (yp, cbp, crp), g = component_planes(rgb)
    ql, qc = quant_tables(quality)
    yq = quantize(fdct_islow(_blocks_of(yp) - 128), ql).reshape(yp.shape[0] // 8, yp.shape[1] // 8, 64)
    cq = [quantize(fdct_islow(_blocks_of(p) - 128), qc).reshape(p.shape[0] // 8, p.shape[1] // 8, 64)
          for p in (cbp, crp)]
    R, C = g["mcu_rows"], g["mcu_cols"]
    out = np.zeros((R, C, 6, 64), dtype=np.int64)
    for r in range(R):
        for c in range(C):
            for k in range(4):                                     # Y0 Y1 / Y2 Y3
                by, bx = 2 * r + (k >> 1), 2 * c + (k & 1)
                if by < g["yb_h"] and bx < g["yb_w"]:
                    out[r, c, k] = yq[by, bx]
                else:                                              # jccoefct.c: dummy block = zero AC, previous DC
                    out[r, c, k, 0] = out[r, c, k - 1, 0] if bx >= g["yb_w"] and by < g["yb_h"] else out[r, c, 1, 0]
            out[r, c, 4] = cq[0][r, c]
            out[r, c, 5] = cq[1][r, c]
    return out.reshape(R * C, 6, 64)[..., ZIGZAG], g


class _BitWriter:
    def __init__(self):
        self.acc, self.n, self.out = 0, 0, bytearray()

    def put(self, code: int, length: int):
        self.acc = (self.acc << length) | (code & ((1 << length) - 1))
        self.n += length
        while self.n >= 8:
            byte = (self.acc >> (self.n - 8)) & 0xFF
            self.out.append(byte)
            if byte == 0xFF:
                self.out.append(0)                                 # byte stuffing
            self.n -= 8
        self.acc &= (1 << self.n) - 1

    def flush(self):
        if self.n:
            self.put(0x7F, 8 - self.n)                             # jchuff.c flush_bits: pad with ones


def entropy_encode(zz: np.ndarray) -> bytes:
    """jchuff.c encode_mcu_huff over [n_mcu, 6, 64] zigzag-ordered quantised blocks."""
    dc = [derive_codes(*DC_LUMA), derive_codes(*DC_CHROMA)]
    ac = [derive_codes(*AC_LUMA), derive_codes(*AC_CHROMA)]
    bw = _BitWriter()
    last_dc = [0, 0, 0]
    comp_of = (0, 0, 0, 0, 1, 2)
    zz_list = zz.tolist()
    for mcu in zz_list:
        for k, blk in enumerate(mcu):
            ci = comp_of[k]
            t = 0 if ci == 0 else 1
            diff = blk[0] - last_dc[ci]
            last_dc[ci] = blk[0]
            mag = -diff if diff < 0 else diff
            nbits = mag.bit_length()
            bw.put(*dc[t][nbits])
            if nbits:
                bw.put(diff if diff >= 0 else diff - 1, nbits)      # negative: low bits of (diff - 1)
            run = 0
            for v in blk[1:]:
                if v == 0:
                    run += 1
                    continue
                while run > 15:
                    bw.put(*ac[t][0xF0])
                    run -= 16
                mag = -v if v < 0 else v
                nbits = mag.bit_length()
                bw.put(*ac[t][(run << 4) | nbits])
                bw.put(v if v >= 0 else v - 1, nbits)
                run = 0
            if run:
                bw.put(*ac[t][0])                                   # EOB
    bw.flush()
    return bytes(bw.out)


def header(height: int, width: int, quality: int = 75) -> bytes:
    """SOI, JFIF APP0, two DQT, SOF0 (4:2:0), four DHT, SOS -- what jpeg_start_compress writes for
    Pillow's defaults (623 bytes)."""
    ql, qc = quant_tables(quality)
    h = bytearray(b"\xff\xd8")
    h += b"\xff\xe0" + (16).to_bytes(2, "big") + b"JFIF\x00\x01\x01\x00\x00\x01\x00\x01\x00\x00"
    for tq, q in ((0, ql), (1, qc)):
        h += b"\xff\xdb" + (67).to_bytes(2, "big") + bytes([tq]) + bytes(int(q[i]) for i in ZIGZAG)
    h += b"\xff\xc0" + (17).to_bytes(2, "big") + b"\x08" + height.to_bytes(2, "big") + width.to_bytes(2, "big")
    h += b"\x03" + bytes([1, 0x22, 0, 2, 0x11, 1, 3, 0x11, 1])
    for tc, (bits, vals) in ((0x00, DC_LUMA), (0x10, AC_LUMA), (0x01, DC_CHROMA), (0x11, AC_CHROMA)):
        h += b"\xff\xc4" + (3 + 16 + len(vals)).to_bytes(2, "big") + bytes([tc]) + bytes(bits) + bytes(vals)
    h += b"\xff\xda" + (12).to_bytes(2, "big") + b"\x03" + bytes([1, 0x00, 2, 0x11, 3, 0x11]) + b"\x00\x3f\x00"
    return bytes(h)


def encode_rgb(rgb: np.ndarray, quality: int = 75) -> bytes:
    """``Image.fromarray(rgb, 'RGB').save(buf, 'JPEG', quality=quality)`` byte for byte."""
    rgb = np.ascontiguousarray(rgb, dtype=np.uint8)
    zz, _ = quantized_blocks(rgb, quality)
    return header(rgb.shape[0], rgb.shape[1], quality) + entropy_encode(zz) + b"\xff\xd9"
