"""NECESSARY DRAM bytes of a fused chain: what any kernel must move at the least, sector by sector.

SURVEY 8d fixes the ALGORITHMIC figure (whole input frame + output). A fused bilinear DOWN-scaler does not need the
whole frame: output row oy blends source rows yi and yi+1 only, output column ox source pixels xi and xi+1 only.
DRAM moves 32-byte sectors, so the floor is

    sum over planes, over SAMPLED rows, of 32 B x (sectors holding a sampled pixel)  +  output bytes.

`roofline.frac` in bench.py is quoted on this figure (it cannot exceed 1 by skipping rows); the algorithmic fraction
is printed beside it. The sampling positions follow libyuv's ScaleSlope for kFilterBilinear exactly as the plan
compiler computes them (band_b200/csrc/band/buffer/scale_dispatch.cc: FixedDiv, centred start when shrinking), so the
figure is derived from the same numbers the kernel runs on. Measurement support only: nothing here computes pixels.
"""
import numpy as np

SECTOR = 32


def _axis(src, dst):
    """(start, step) in 16.16 along a filtered axis; centred when shrinking, edge to edge when enlarging."""
    if dst <= src:
        step = (src << 16) // dst
        return (step >> 1) - 32768, step
    if dst > 1:
        return 0, ((src << 16) - 0x00010001) // (dst - 1)
    return 0, 0


def sampled(src, dst):
    """Sorted unique source indices the bilinear filter reads along one axis (position i and i+1 when the fraction
    is not zero), clamped like the kernels clamp (max index src-1)."""
    start, step = _axis(src, dst)
    pos = start + np.arange(dst, dtype=np.int64) * step
    pos = np.clip(pos, 0, (src - 1) << 16)
    i = pos >> 16
    frac = (pos >> 8) & 255
    second = np.where(frac != 0, np.minimum(i + 1, src - 1), i)
    return np.unique(np.concatenate([i, second]))


def _sectors(cols, bytes_per_px, row_bytes):
    """32-byte sectors of one row touched by the pixels `cols` (rows start sector aligned at these sizes)."""
    lo = cols * bytes_per_px
    hi = lo + bytes_per_px - 1
    s = np.unique(np.concatenate([lo // SECTOR, hi // SECTOR]))
    return int(s.size), (row_bytes + SECTOR - 1) // SECTOR


def necessary_yuv420_to_rgb(src_w, src_h, dst_w, dst_h, out_bytes_per_px, interleaved=True,
                            region=None, swap_axes=False):
    """4:2:0 source -> RGB bilinear resize. region = (x0, y0, w, h) of the source the chain looks at (Crop, in source
    coordinates); swap_axes: a quarter-turn sits in front, so output rows walk source columns. Returns a dict with the
    necessary bytes and their split."""
    x0, y0, w, h = region or (0, 0, src_w, src_h)
    # the resize sees a w x h image (h x w after a quarter-turn); along source rows it samples `rows`, along columns `cols`
    if swap_axes:
        rows = sampled(h, dst_w) + y0
        cols = sampled(w, dst_h) + x0
    else:
        rows = sampled(h, dst_h) + y0
        cols = sampled(w, dst_w) + x0
    n_sec, _ = _sectors(cols, 1, src_w)
    luma = int(rows.size) * n_sec * SECTOR
    crows = np.unique(rows >> 1)
    ccols = np.unique(cols >> 1)
    if interleaved:
        c_sec, _ = _sectors(ccols, 2, 2 * ((src_w + 1) // 2))
        chroma = int(crows.size) * c_sec * SECTOR
    else:
        c_sec, _ = _sectors(ccols, 1, (src_w + 1) // 2)
        chroma = 2 * int(crows.size) * c_sec * SECTOR
    out = dst_w * dst_h * out_bytes_per_px
    return {"necessary_bytes": luma + chroma + out, "luma_rows": int(rows.size), "chroma_rows": int(crows.size),
            "luma_bytes": luma, "chroma_bytes": chroma, "out_bytes": out}


def necessary_packed(src, dst, src_bpp, out_bytes_per_px):
    """Square packed source (gray / RGB / RGBA, src_bpp bytes per pixel) -> dst x dst. Up-scales and same-size
    conversions read every source byte; down-scales only the sampled sectors of the sampled rows."""
    out = dst * dst * out_bytes_per_px
    if dst >= src:
        return {"necessary_bytes": src * src * src_bpp + out, "out_bytes": out}
    idx = sampled(src, dst)
    # exact even integer ratios take libyuv's 2x2 box at the sample origin: rows/cols i and i+1 of every step
    step = (src << 16) // dst
    if step & 0xffff == 0 and (step >> 16) % 2 == 0:
        k = step >> 16
        base = np.arange(dst, dtype=np.int64) * k + (k >> 1) - 1
        idx = np.unique(np.concatenate([base, base + 1]))
    n_sec, _ = _sectors(idx, src_bpp, src * src_bpp)
    return {"necessary_bytes": int(idx.size) * n_sec * SECTOR + out, "rows": int(idx.size), "out_bytes": out}
