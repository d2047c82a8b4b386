// tile_map.cuh -- output-tile raster of the persistent tcgen05 GEMM kernels (host- and device-callable so the range maths is
// unit-tested on the CPU through ckb_tile_coord, in the style of the reference's test/block_to_ctile_map).
#pragma once

namespace ckb {

#ifdef __CUDACC__
#define CKB_HD __host__ __device__ __forceinline__
#else
#define CKB_HD inline
#endif

CKB_HD void tile_coord(int t, int num_m_blk, int num_n_blk, int& mb, int& nb, int GROUP = 8)
{
    // bands of GROUP M-blocks; inside a band N-major so concurrently running CTAs share B panels and
    // a band of A stays L2 resident (the role of CK's BlockToCTileMap_Grouped_M00_N0_M01Adapt,
    // grid/block_to_ctile_map.hpp:266-390).
    const int per_band  = GROUP * num_n_blk;
    const int band      = t / per_band;
    const int first_m   = band * GROUP;
    const int rows      = (GROUP < num_m_blk - first_m) ? GROUP : (num_m_blk - first_m);
    const int r         = t - band * per_band;
    mb                  = first_m + r % rows;
    nb                  = r / rows;
}

// stream-K: CTA `cta` of `grid` owns the contiguous range [begin, end) of the flattened (n-tile, k-unit) space -- equal shares,
// the remainder spread over the CTAs (the role of BlockToCTileMap_GemmStreamK, grid/block_to_ctile_map.hpp:1016-1400)
CKB_HD void streamk_range(long long total_units, int cta, int grid, long long& begin, long long& end)
{
    begin = total_units * cta / grid;
    end   = total_units * (cta + 1) / grid;
}

} // namespace ckb
