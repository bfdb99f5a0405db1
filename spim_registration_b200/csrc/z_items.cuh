// Work-item order of the z pass, shared by the plain and the TMA kernel.
#pragma once

namespace mvd {

// Work item -> (position block, tile).  Tiles run fastest so that the CTAs in flight read the same part of
// the kernel spectrum; with MVD_Z_PGROUP = G > 1, G neighbouring position blocks (G * PW * 8 contiguous
// bytes per z index) run fastest of all, so that CTAs in flight also share DRAM pages.
#ifndef MVD_Z_PGROUP
#define MVD_Z_PGROUP 4  // measured (cfg3 class, 4 convolutions): 1 -> 11.14 ms, 4 -> 10.88, 8 -> 10.93, 16 -> 11.03
#endif
__device__ __forceinline__ void decode_item(int item, int n_tiles, int& pblk, int& ti) {
    if constexpr (MVD_Z_PGROUP > 1) {
        const int per_group = MVD_Z_PGROUP * n_tiles;
        const int g = item / per_group, r = item - g * per_group;
        ti = r / MVD_Z_PGROUP;
        pblk = g * MVD_Z_PGROUP + (r - ti * MVD_Z_PGROUP);
    } else {
        pblk = item / n_tiles;
        ti = item - pblk * n_tiles;
    }
}

}  // namespace mvd
