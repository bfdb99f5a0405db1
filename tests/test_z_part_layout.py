"""The shared-memory layout of z_part_kernel (csrc/z_warp_kernels.cu), restated on the host: lane -> (line, lane of the
line), landing rows, exchange rows (xrow) and the per-lane offset table.  Checks what the kernel's design rests on:

  * the exchange is a bijection inside the warp's own 16-byte column of the swizzled item buffer (a warp never touches
    another warp's data, so no barrier between warps is needed);
  * what a lane writes as output r of its radix-32 butterfly (element 32 l + r) is what the lane that needs element
    l' + LPL m reads;
  * every 8-byte access of a half-warp touches 16 different bank pairs (2 wavefronts per warp instruction, the minimum).

No GPU needed; the device code is checked against the oracle in tests/test_gpu_parity.py (z_warp = 3 / 4).
"""
import pytest


def swz(z, p):
    """zw_swz: byte offset of element (row z, position p) in a buffer written by TMA with the 128-byte swizzle"""
    return z * 128 + (((p >> 1) ^ (z & 7)) << 4) + (p & 1) * 8


def xrow(i):
    return i ^ ((i >> 5) & 7)


def lanes_of_warp(bz, w):
    lpl = bz // 32
    lpw = 32 // lpl
    out = []
    for t in range(32):
        if lpl == 16:
            h, l = (t >> 3) & 1, (t & 7) | ((t >> 4) << 3)
        else:
            h, l = t >> 3, t & 7
        p = lpw * w + h
        pc, hb, l7 = p >> 1, (p & 1) * 8, l & 7
        xo = [((c ^ l7) << 7) + (((pc ^ l7 ^ c) & 7) << 4) + hb for c in range(8)]
        out.append(dict(t=t, h=h, l=l, p=p, xo=xo, put_base=l * 4096, get_base=(l & 8) * 128))
    return out


def wavefronts(addrs):
    """8-byte accesses are served per half-warp; a wavefront serves one access per bank pair"""
    n = 0
    for half in (addrs[:16], addrs[16:]):
        banks = {}
        for a in half:
            banks.setdefault((a // 8) % 16, set()).add(a)
        n += max(len(v) for v in banks.values())
    return n


@pytest.mark.parametrize("bz", [512, 256])
def test_z_part_exchange_is_a_conflict_free_bijection_inside_the_warps_column(bz):
    lpl = bz // 32
    lpw = 32 // lpl
    nwarp = 16 // lpw
    seen_positions = set()
    for w in range(nwarp):
        lanes = lanes_of_warp(bz, w)
        put, get, land = {}, {}, {}
        for ln in lanes:
            seen_positions.add(ln["p"])
            for r in range(32):
                put[(ln["h"], 32 * ln["l"] + r)] = ln["put_base"] + ln["xo"][r & 7] + 1024 * (r >> 3)
            for m in range(32):
                c = (lpl * m) >> 5
                get[(ln["h"], ln["l"] + lpl * m)] = ln["get_base"] + ln["xo"][c & 7] + 128 * ((lpl * m) & ~7)
                land[(ln["h"], ln["l"] + lpl * m)] = swz(ln["l"], ln["p"]) + lpl * 128 * m
        assert put == get                                        # the reader finds what the writer left
        assert len(put) == lpw * bz and len(set(put.values())) == len(put)
        for (h, i), a in put.items():                            # row xrow(i) of the line's own column
            assert a == swz(xrow(i), lpw * w + h)
            assert 0 <= a < 16 * bz * 8
        for (h, z), a in land.items():                           # the item's data: row z of the same column
            assert a == swz(z, lpw * w + h)
        for r in range(32):
            assert wavefronts([ln["put_base"] + ln["xo"][r & 7] + 1024 * (r >> 3) for ln in lanes]) == 2
        for m in range(32):
            c = (lpl * m) >> 5
            assert wavefronts([ln["get_base"] + ln["xo"][c & 7] + 128 * ((lpl * m) & ~7) for ln in lanes]) == 2
            assert wavefronts([swz(ln["l"], ln["p"]) + lpl * 128 * m for ln in lanes]) == 2
    assert seen_positions == set(range(16))


def test_xrow_permutes_every_block_of_32_rows():
    for bz in (256, 512):
        rows = [xrow(i) for i in range(bz)]
        assert sorted(rows) == list(range(bz))
        assert all(r >> 5 == i >> 5 for i, r in enumerate(rows))   # only the low three bits move


def test_first_generation_layout_had_four_wavefronts_per_access():
    """z_warp_kernel: one position per warp, lane t holds rows t + 32 m -- 8 chunks x one half: 4 wavefronts (why the
    second generation exists)."""
    for w in range(16):
        assert wavefronts([swz(t, w) for t in range(32)]) == 4
