// state.cuh — layout of the three opaque scratch buffers (geometry / binning / image state).
// Plays the role of GeometryState/BinningState/ImageState::fromChunk (HR/cuda_rasterizer/rasterizer_impl.cu:159-199)
// but with its own SoA layout: clamp flags packed to one byte, rects kept inside, the scan's look-back
// states and one 48-byte blend record per Gaussian added, no CUB temp.
#pragma once
#include "common.cuh"

namespace lodgs {

constexpr int PRE_BLOCK = 256;  // Gaussians per preprocess tile (one look-back state each)

constexpr int SORT_RADIX = 256;
constexpr int SORT_MAX_PASSES = 8;
constexpr int SORT_TILE = 4096;  // keys per onesweep tile (256 threads x 16 keys)
static inline size_t sort_tiles(size_t n) { return (n + SORT_TILE - 1) / SORT_TILE; }
// histograms [8][256] u32 + tickets [8] i32 + look-back [8][tiles][256] u32, each 128-byte aligned (see sort_from_chunk)
static inline size_t sort_ws_bytes(size_t n) {
    return align_up(SORT_MAX_PASSES * SORT_RADIX * 4, 128) + align_up(SORT_MAX_PASSES * 4, 128) +
           align_up((size_t)SORT_MAX_PASSES * sort_tiles(n) * SORT_RADIX * 4, 128);
}

// Blend inputs of one Gaussian, written by the preprocess kernel and gathered into shared memory by the blend kernels
// through the tile's sorted id list (three 16-byte async copies per record).
struct __align__(16) SplatRec {
    float x, y, conx, cony;      // projected mean, conic xx, xy
    float conz, opac, r, g;      // conic yy, opacity (AA-compensated), colour
    float b, invdepth, t, invkids;  // colour, 1/view depth (rcp.rn), LOD weight (alpha cut-off without LOD), 1/kids
};
static_assert(sizeof(SplatRec) == 48, "SplatRec must be 48 bytes");

struct GeomHeader {
    unsigned num_rendered;  // total (Gaussian,tile) instances, written by the last preprocess tile
    int ticket;             // dynamic tile dispenser for the fused scan
    int ticket2;            // same for the depth-order scan
    int n_visible;          // entries with a depth key (not culled), written by depth_compact_kernel when the compacted depth sort is used
    int pad[28];
};

struct GeomState {
    GeomHeader* hdr;
    unsigned long long* scan_state;  // [ceil(P/PRE_BLOCK)]
    SplatRec* recs;                  // [P] blend inputs of every visible Gaussian (gathered by the blend kernels)
    float* cov3D;                    // [P,6]
    uint8_t* clamped;                // [P] bit c set = channel c clamped to 0
    uint32_t* tiles_touched;         // [P]
    int2* rects;                     // [P]
    // depth-ordered emission (raster_api.cu): Gaussians sorted by depth bits, then their tile counts scanned in that order
    uint32_t* depth_keys;            // [P] depth bits, 0xffffffff for culled Gaussians
    uint32_t* depth_keys_sorted;     // [P]
    uint32_t* order;                 // [P] list positions in (depth bits, index) order
    uint32_t* sort_tmp_k;            // [P] ping-pong
    uint32_t* sort_tmp_v;            // [P]
    uint32_t* sorted_offsets;        // [P] inclusive scan of the tile counts in depth order
    uint2* tile_box;                 // [P] tile rectangle {min.x | min.y<<16, max.x | max.y<<16}, 0 for culled Gaussians
    // render-list mode: {depth key, list position} of the visible entries in list order (binning.cu depth_compact_kernel). They are the INPUT
    // of the depth sort's first pass only, and that pass (of four) writes the ping-pong pair, so they live in the sort's OUTPUT arrays.
    uint32_t* compact_keys;          // = depth_keys_sorted
    uint32_t* compact_vals;          // = order
    unsigned long long* scan2_state; // [ceil(P/PRE_BLOCK)+1] look-back states of the emission's scan
    void* sort_ws;                   // radix sort workspace for P pairs
    size_t offsets[9];
    size_t bytes;
    size_t zero_bytes;  // header + scan states: cleared at the start of every forward
};

template <typename T>
static inline T* carve_arr(char* base, size_t& off, size_t count) {
    off = align_up(off, 128);
    T* p = reinterpret_cast<T*>(base + off);
    off += count * sizeof(T);
    return p;
}

static inline GeomState geom_from_chunk(void* chunk, size_t P) {
    GeomState g;
    char* b = (char*)chunk;
    size_t off = 0;
    g.hdr = carve_arr<GeomHeader>(b, off, 1);
    g.scan_state = carve_arr<unsigned long long>(b, off, (P + PRE_BLOCK - 1) / PRE_BLOCK + 1);
    g.scan2_state = carve_arr<unsigned long long>(b, off, (P + PRE_BLOCK - 1) / PRE_BLOCK + 1);
    g.zero_bytes = off;
    g.depth_keys = carve_arr<uint32_t>(b, off, P);     g.offsets[0] = (char*)g.depth_keys - b;
    g.recs = carve_arr<SplatRec>(b, off, P);           g.offsets[1] = g.offsets[2] = g.offsets[3] = (char*)g.recs - b;
    g.cov3D = carve_arr<float>(b, off, P * 6);         g.offsets[4] = (char*)g.cov3D - b;
    g.clamped = carve_arr<uint8_t>(b, off, P);         g.offsets[5] = (char*)g.clamped - b;
    g.tiles_touched = carve_arr<uint32_t>(b, off, P);  g.offsets[6] = (char*)g.tiles_touched - b;
    g.rects = carve_arr<int2>(b, off, P);              g.offsets[7] = g.offsets[8] = (char*)g.rects - b;
    g.depth_keys_sorted = carve_arr<uint32_t>(b, off, P);
    g.order = carve_arr<uint32_t>(b, off, P);
    g.sort_tmp_k = carve_arr<uint32_t>(b, off, P);
    g.sort_tmp_v = carve_arr<uint32_t>(b, off, P);
    g.sorted_offsets = carve_arr<uint32_t>(b, off, P);
    g.tile_box = carve_arr<uint2>(b, off, P);
    g.compact_keys = g.depth_keys_sorted;
    g.compact_vals = g.order;
    off = align_up(off, 128);
    g.sort_ws = b + off;
    off += sort_ws_bytes(P);
    g.bytes = align_up(off, 128);
    return g;
}

struct SortWs {
    unsigned* hist;                  // [SORT_MAX_PASSES][256] global digit histograms -> exclusive bases
    int* tickets;                    // [SORT_MAX_PASSES]
    unsigned* lookback;              // [passes][tiles][256] packed {flag:2, count:30}
    size_t bytes;
    size_t zero_bytes;
};
static inline SortWs sort_from_chunk(void* chunk, size_t n) {
    SortWs w;
    char* b = (char*)chunk;
    size_t off = 0;
    w.hist = carve_arr<unsigned>(b, off, SORT_MAX_PASSES * SORT_RADIX);
    w.tickets = carve_arr<int>(b, off, SORT_MAX_PASSES);
    w.lookback = carve_arr<unsigned>(b, off, (size_t)SORT_MAX_PASSES * sort_tiles(n) * SORT_RADIX);
    w.bytes = align_up(off, 128);
    w.zero_bytes = w.bytes;
    return w;
}

struct BinningState {
    uint32_t* tiles_unsorted;  // [R] tile id of every instance, emitted in (depth, index) order of the Gaussians
    uint32_t* vals_unsorted;   // [R] list position of the instance's Gaussian, same order
    uint32_t* tiles_sorted;    // [R] tile ids after the stable sort by tile
    uint32_t* point_list;      // [R] list positions of the Gaussians, sorted by (tile, depth bits, index)
    void* sort_ws;
    size_t offsets[5];
    size_t bytes;
};
static inline BinningState binning_from_chunk(void* chunk, size_t R) {
    BinningState s;
    char* b = (char*)chunk;
    size_t off = 0;
    s.tiles_unsorted = carve_arr<uint32_t>(b, off, R); s.offsets[0] = (char*)s.tiles_unsorted - b;
    s.vals_unsorted = carve_arr<uint32_t>(b, off, R);  s.offsets[1] = (char*)s.vals_unsorted - b;
    s.tiles_sorted = carve_arr<uint32_t>(b, off, R);   s.offsets[2] = (char*)s.tiles_sorted - b;
    s.point_list = carve_arr<uint32_t>(b, off, R + 8); s.offsets[3] = s.offsets[4] = (char*)s.point_list - b;
    off = align_up(off, 128);
    s.sort_ws = b + off;
    off += sort_from_chunk(nullptr, R).bytes;
    s.bytes = align_up(off, 128);
    return s;
}

struct ImageState {
    float* final_T;       // [W*H]
    uint32_t* n_contrib;  // [W*H]
    uint2* ranges;        // [tiles]
    uint32_t* tile_order; // [tiles] tile ids, longest list first: the order the blend CTAs take them in; [tiles] = number of LONG tiles (the first ones)
    size_t offsets[3];
    size_t bytes;
};
static inline ImageState image_from_chunk(void* chunk, int W, int H) {
    ImageState s;
    char* b = (char*)chunk;
    size_t off = 0;
    size_t N = (size_t)W * H;
    size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
    s.final_T = carve_arr<float>(b, off, N);        s.offsets[0] = (char*)s.final_T - b;
    s.n_contrib = carve_arr<uint32_t>(b, off, N);   s.offsets[1] = (char*)s.n_contrib - b;
    s.ranges = carve_arr<uint2>(b, off, tiles);     s.offsets[2] = (char*)s.ranges - b;
    s.tile_order = carve_arr<uint32_t>(b, off, tiles + 4);
    s.bytes = align_up(off, 128);
    return s;
}

}  // namespace lodgs
