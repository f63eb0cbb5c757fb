// Persistent variant of the fused level kernel for levels > 0 in batch mode (see pyramid.cuh for the tile phases).
//
// What changes against pyr_level_kernel<false, 32>:
//   * the source box of a tile (rows [sy_lo, ..] x 16-byte aligned columns [sx_lo, ..] of level l-1's bordered plane)
//     arrives by ONE bulk tensor copy (TMA, cp.async.bulk.tensor.3d: x = byte in the plane row, y = plane row,
//     z = frame) signalled on an mbarrier, instead of 256 threads' LDG -> STS loops with their address arithmetic;
//     columns/rows of the box past the plane are zero-filled by the copy engine and never read (the reflect of
//     REFLECT_101 is in the tap tables, not in the staged pixels);
//   * a CTA is persistent (grid = SMs x resident CTAs) and walks tiles in a fixed stride; the box of tile t+1 is
//     requested while tile t is in its FAST phase, so the copy latency is off the critical path;
//   * the vertical blur runs before FAST: after it, the T and Hs areas are dead and receive the next tile's box
//     and tap tables, so the prefetch costs no shared memory (54 KB per CTA, four CTAs per SM, as before).
//
// Shared memory (TROWS = 38):   region A = [0, 15200)          region B = [15200, 54112)      mbarrier at 54112
//   early (building T):          S box | .. | tap tables        Hq (horizontally resized rows)
//   late  (consuming T):         T | Hs                         pair planes PA, PB
#pragma once
#include <cuda.h>
#include "pyramid.cuh"

namespace orbx {

constexpr int PYRT_TH = TH_BATCH;
constexpr int PYRT_TROWS = PYRT_TH + 2 * HALO;
__host__ __device__ constexpr size_t pyrt_region_a() { return (size_t)PYRT_TROWS * TS + (size_t)PYRT_TROWS * TW * 2; }
__host__ __device__ constexpr size_t pyrt_region_b() { return (size_t)2 * PYRT_TROWS * 32 * sizeof(uint4); }
__host__ __device__ constexpr size_t pyrt_taps_off() { return pyrt_region_a() - ((size_t)TBYTES * sizeof(TapX) + 40 * sizeof(TapX)); }
__host__ __device__ constexpr size_t pyrt_smem_bytes() { return pyrt_region_a() + pyrt_region_b() + 16; }
// the level can take this kernel: the source box fits under the tap tables in region A, Hq fits in region B,
// and the box obeys the tensor-map limits (each box dimension <= 256, inner dimension a multiple of 16 bytes)
__host__ inline bool pyrt_fits(int srcRows, int srcCols) {
    return srcCols % 16 == 0 && srcCols <= 256 && srcRows <= 256 && (size_t)srcRows * srcCols <= pyrt_taps_off() &&
           pyr_smem_Hq(srcRows) <= pyrt_region_b();
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}
// 3-D tiled bulk tensor load global -> shared, completion counted in bytes on `bar`
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"((unsigned long long)map), "r"(c0), "r"(c1), "r"(c2), "r"(bar) : "memory");
}

struct TileId { int frame, tx, ty; };

__global__ void __launch_bounds__(PYR_THREADS, 4) pyr_level_tma_kernel(const __grid_constant__ PyrArgs a,
                                                                       const __grid_constant__ CUtensorMap srcMap,
                                                                       int tilesX, int tilesY, int frames) {
    constexpr int TH = PYRT_TH, TROWS = PYRT_TROWS;
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* const S = smem;                                                     // early, region A
    TapX* const tapx = (TapX*)(smem + pyrt_taps_off());
    TapX* const tapy = tapx + TBYTES;
    unsigned short* const Hq = (unsigned short*)(smem + pyrt_region_a());        // early, region B
    uint8_t* const T = smem;                                                     // late, region A
    unsigned short* const Hs = (unsigned short*)(smem + (size_t)TROWS * TS);
    uint4* const PA = (uint4*)(smem + pyrt_region_a());                          // late, region B
    uint4* const PB = PA + TROWS * 32;
    const uint32_t bar = smem_u32(smem + pyrt_region_a() + pyrt_region_b());
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int perFrame = tilesX * tilesY, total = perFrame * frames;
    const uint32_t boxBytes = (uint32_t)(a.srcTileRows * a.srcTileCols);

    auto decode = [&](int t) {
        TileId id;
        id.frame = t / perFrame;
        const int r = t - id.frame * perFrame;
        id.ty = r / tilesX; id.tx = r - id.ty * tilesX;
        return id;
    };
    // the stride between a CTA's consecutive tiles, decomposed once: advancing is three adds with carries, no division
    const TileId step = decode((int)gridDim.x);
    auto advance = [&](TileId id) {
        id.tx += step.tx; id.ty += step.ty; id.frame += step.frame;
        if (id.tx >= tilesX) { id.tx -= tilesX; id.ty++; }
        if (id.ty >= tilesY) { id.ty -= tilesY; id.frame++; }
        return id;
    };
    auto request = [&](const TileId& id) {               // one thread: arm the barrier, start the copy of the tile's source box
        const int sx_lo = (int)(a.tileX[id.tx] & 0xffffu), sy_lo = (int)(a.tileY[id.ty] & 0xffffu);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // region A was last written through the generic proxy
        mbar_expect_tx(bar, boxBytes);
        tma_load_3d(smem_u32(S), &srcMap, a.srcX0 + sx_lo, ORB_EDGE + sy_lo, id.frame, bar);
    };

    int t = blockIdx.x;
    if (t >= total) return;
    TileId cur = decode(t);
    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        request(cur);
    }
    tile_build_taps<TH>(a, tapx, tapy, cur.tx * TW, cur.ty * TH, (int)(a.tileX[cur.tx] & 0xffffu), (int)(a.tileY[cur.ty] & 0xffffu), tid);
    __syncthreads();                                     // barrier initialised, taps of the first tile complete
    uint32_t parity = 0;
    for (;;) {
        const int x0 = cur.tx * TW, y0 = cur.ty * TH, frame = cur.frame;
        const uint32_t ex = a.tileX[cur.tx], ey = a.tileY[cur.ty];
        const int nrows = (int)(ey >> 16) - (int)(ey & 0xffffu) + 1;
        mbar_wait(bar, parity);                          // the box has landed (async proxy writes made visible by the barrier)
        parity ^= 1;
        tile_hresize(a, S, a.srcTileCols, nrows, tapx, Hq, x0, tid);
        __syncthreads();                                 // S and tapx consumed, Hq complete
        tile_vresize<TH>(a, tapy, Hq, T, x0, y0, tid);   // T overwrites the box
        __syncthreads();                                 // T complete, Hq and tapy consumed
        tile_store_plane<TH>(a, T, frame, x0, y0, tid);
        tile_row_pass<TH>(T, Hs, PA, PB, a.h, y0, wid, lane);
        __syncthreads();                                 // planes and Hs complete, T consumed
        tile_vblur_pass<TH>(a, Hs, frame, x0, y0, wid, lane);
        __syncthreads();                                 // region A is dead: it takes the next tile's box and tap tables
        const int tn = t + (int)gridDim.x;
        const bool more = tn < total;
        TileId nxt = cur;
        if (more) {
            nxt = advance(cur);
            if (tid == 0) request(nxt);
        }
        tile_fast_pass<TH>(a, PA, PB, frame, x0, y0, wid, lane);
        if (!more) break;
        tile_build_taps<TH>(a, tapx, tapy, nxt.tx * TW, nxt.ty * TH, (int)(a.tileX[nxt.tx] & 0xffffu), (int)(a.tileY[nxt.ty] & 0xffffu), tid);
        __syncthreads();                                 // planes consumed (region B takes Hq), taps of the next tile complete
        t = tn; cur = nxt;
    }
}

}  // namespace orbx
