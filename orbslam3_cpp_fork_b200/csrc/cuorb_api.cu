// libcuorb — C ABI implementation (see include/cuorb.h): extractor handle, geometry tables
// computed with the reference's float arithmetic, stream-ordered batch pipeline.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "common.cuh"
#include "tma.cuh"

namespace cuorb {

static thread_local std::string t_last_error;
thread_local bool t_pdl = false;   // common.cuh: launch the extraction kernels as programmatic dependents
void set_error(const std::string& msg) { t_last_error = msg; }
std::atomic<unsigned long long> g_launch_count{0};

static inline int cv_round_f(float v) { return (int)lrintf(v); }
static inline short sat_short(float v) {
  int i = cv_round_f(v);
  return (short)std::min(std::max(i, -32768), 32767);
}
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace cuorb

using namespace cuorb;

struct cuorb_extractor {
  cuorb_config  cfg;
  ExtractorGeom g;  // geometry of the CURRENT image size (rebuilt when the size changes)
  int           cur_w = 0, cur_h = 0;
  float scale[MAX_LEVELS], inv_scale[MAX_LEVELS], sigma2[MAX_LEVELS], inv_sigma2[MAX_LEVELS];
  int   quota[MAX_LEVELS];

  cudaStream_t stream = nullptr;
  cudaStream_t side_stream = nullptr;   // blur runs here, beside FAST + quadtree (both only read the pyramid)
  cudaEvent_t  ev_done = nullptr;       // recorded after the last enqueued work
  cudaEvent_t  ev_pyr = nullptr, ev_blur = nullptr;   // fork / join of the side stream
  cudaEvent_t  ev_stage[CUORB_STAGE_COUNT + 1] = {};
  bool         profiling = false, have_stage_times = false, stereo_timed = false;

  // device buffers (sized for max_width x max_height x max_batch)
  ExtractorGeom* d_geom = nullptr;
  ResizeTap *d_xtab = nullptr, *d_ytab = nullptr;
  uint4*     d_wtab = nullptr;        // per-word horizontal tap tables of k_pyr_resize_stage
  int        wtab_off[MAX_LEVELS];
  PyrPlan    pyr_plan[MAX_LEVELS];    // staged resize kernel: CTA shape and staged source window per level
  size_t     cap_wtab = 0;
  int4*      d_blur_tiles = nullptr;
  TmaMaps    tmaps;                   // TMA tensor maps of the pyramid / blurred planes (per level), passed as __grid_constant__
  int4*      d_fast_cells = nullptr;  // per FAST cell: level, x0 | y0 << 16, dw | dh << 16
  int4*      d_fast_recs = nullptr;   // k_fast_persist work records (2 per non-empty cell)
  int        n_fast_recs = 0;
  bool       fast_persist = false;
  FastMaps   fmaps;
  int        n_blur_tiles = 0;
  uint8_t *  d_input = nullptr, *d_pyr = nullptr, *d_blur = nullptr;
  unsigned * d_cand = nullptr, *d_cand_scratch = nullptr;
  int *      d_cand_count = nullptr, *d_key_count = nullptr, *d_flags = nullptr, *d_slot = nullptr;
  LevelKey*  d_keys = nullptr;
  cuorb_keypoint* d_out_kps = nullptr;
  uint8_t*   d_out_desc = nullptr;
  int *      d_out_n = nullptr, *d_out_mono = nullptr;
  float *    d_uright = nullptr, *d_depth = nullptr;
  int *      d_sad = nullptr, *d_best_idx = nullptr, *d_best_dist = nullptr;
  uint4*     d_rtab = nullptr;        // stereo: right keypoints bucketed by row (per image kp_cap entries)
  int*       d_rstart = nullptr;      // stereo: bucket starts (per image STEREO_MAX_BUCKETS + 1)
  // pinned host mirrors of the small per-image results
  int *h_out_n = nullptr, *h_out_mono = nullptr, *h_flags = nullptr;
  size_t cap_input = 0, cap_pyr = 0, cap_blur = 0, cap_cand = 0, cap_keys = 0, cap_kp = 0;
  size_t cap_xtab = 0, cap_ytab = 0, cap_tiles = 0, cap_cells = 0;
  int    n_img = 0;       // images in the last batch
  int    fetch_n = 0, fetch_capacity = 0;   // last fetch: images copied to h_out_n / caller capacity (0: no kps/desc asked)
  // d_flags, d_out_n, d_out_mono, d_out_kps, d_out_desc, d_uright, d_depth are views into ONE allocation (d_pack), in that order:
  // a handle made for one image at a time (max_batch == 1, what the SLAM drop-in creates) fetches everything a
  // frame produces with a single D2H into the pinned mirror h_pack
  uint8_t *d_pack = nullptr, *h_pack = nullptr;
  size_t   pack_bytes = 0, pack_n_off = 0, pack_mono_off = 0, pack_kps_off = 0, pack_desc_off = 0, pack_ur_off = 0, pack_depth_off = 0;
  bool   have_batch = false, have_stereo = false;
};

namespace cuorb {

// ctor tables (src/ORBextractor.cc:411-440)
static void build_tables(cuorb_extractor* e) {
  const cuorb_config& c = e->cfg;
  const double scaleFactor = (double)c.scale_factor;  // float ctor arg -> double member
  e->scale[0] = 1.0f; e->sigma2[0] = 1.0f;
  for (int i = 1; i < c.nlevels; i++) {
    e->scale[i]  = (float)(e->scale[i - 1] * scaleFactor);
    e->sigma2[i] = e->scale[i] * e->scale[i];
  }
  for (int i = 0; i < c.nlevels; i++) {
    e->inv_scale[i]  = 1.0f / e->scale[i];
    e->inv_sigma2[i] = 1.0f / e->sigma2[i];
  }
  float factor = (float)(1.0f / scaleFactor);
  float nDesired = (float)(c.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)c.nlevels)));
  int sum = 0;
  for (int l = 0; l < c.nlevels - 1; l++) {
    e->quota[l] = cv_round_f(nDesired);
    sum += e->quota[l];
    nDesired *= factor;
  }
  e->quota[c.nlevels - 1] = std::max(c.nfeatures - sum, 0);
}

// Per-size geometry.  Returns CUORB_OK or an error for sizes the reference has UB for.
static int build_geometry(const cuorb_extractor* e, int w, int h, ExtractorGeom& g,
                          std::vector<ResizeTap>& xtab, std::vector<ResizeTap>& ytab) {
  const cuorb_config& c = e->cfg;
  memset(&g, 0, sizeof(g));
  g.nlevels = c.nlevels; g.ini_th = c.ini_th_fast; g.min_th = c.min_th_fast;
  xtab.clear(); ytab.clear();
  size_t pyr_off = 0, blur_off = 0, cand_off = 0;
  int key_off = 0, cell_base = 0;
  for (int l = 0; l < c.nlevels; l++) {
    LevelGeom& L = g.lv[l];
    const float s = e->inv_scale[l];
    L.w = cv_round_f((float)w * s);  // :1143
    L.h = cv_round_f((float)h * s);
    L.scale = e->scale[l]; L.inv_scale = s;
    L.kp_size = (float)(int)(PATCH_SIZE * e->scale[l]);  // :865
    if (L.w <= 2 * EDGE_THRESHOLD || L.h <= 2 * EDGE_THRESHOLD) {
      set_error("pyramid level too small for the 19-px border / FAST region");
      return CUORB_E_UNSUPPORTED;
    }
    L.pitch     = (int)align_up((size_t)ROW_LEAD + L.w + EDGE_THRESHOLD, 64);
    L.plane_off = pyr_off;
    L.pix_off   = pyr_off + (size_t)EDGE_THRESHOLD * L.pitch + ROW_LEAD;
    pyr_off += align_up((size_t)L.pitch * (L.h + 2 * EDGE_THRESHOLD), 256);
    L.bpitch   = (int)align_up((size_t)L.w, 16);
    L.blur_off = blur_off;
    blur_off += align_up((size_t)L.bpitch * L.h, 256);
    // FAST cells (:761-775)
    L.maxBorderX = L.w - EDGE_THRESHOLD + 3;
    L.maxBorderY = L.h - EDGE_THRESHOLD + 3;
    const float width = (float)(L.maxBorderX - FAST_BORDER), height = (float)(L.maxBorderY - FAST_BORDER);
    L.nCols = (int)(width / 35.f);
    L.nRows = (int)(height / 35.f);
    if (L.nCols <= 0 || L.nRows <= 0) {
      set_error("pyramid level smaller than one 35-px FAST cell (division by zero in the reference)");
      return CUORB_E_UNSUPPORTED;
    }
    L.wCell = (int)std::ceil(width / L.nCols);
    L.hCell = (int)std::ceil(height / L.nRows);
    L.cell_base = cell_base;
    cell_base += L.nCols * L.nRows;
    g.max_cell_w = std::max(g.max_cell_w, L.wCell);
    g.max_cell_h = std::max(g.max_cell_h, L.hCell);
    // octree roots (:566-568)
    L.quota = e->quota[l];
    L.nIni  = (int)std::round(width / height);
    if (L.nIni <= 0) {
      set_error("portrait image with aspect < 0.5: nIni == 0 (division by zero in the reference)");
      return CUORB_E_UNSUPPORTED;
    }
    L.hX       = width / L.nIni;
    L.node_cap = std::max(L.quota, 4 * L.nIni) + 8;
    L.key_cap  = L.node_cap;
    L.key_off  = key_off;
    key_off += L.key_cap;
    // worst case of strict 8-neighbour local maxima: one per 2x2 block of each cell
    L.cand_cap = L.nCols * L.nRows * ((L.wCell + 1) / 2) * ((L.hCell + 1) / 2);
    L.cand_off = cand_off;
    cand_off += align_up((size_t)L.cand_cap, 32);
    if (L.w + 2 * EDGE_THRESHOLD >= 4096 || L.h + 2 * EDGE_THRESHOLD >= 4096 || L.node_cap >= 16384) {
      set_error("image or feature count beyond the packed 12-bit coordinate / 14-bit node range");
      return CUORB_E_UNSUPPORTED;
    }
    // resize tables (OpenCV resize.cpp: scale = 1/(dst/src) in double, float fx, 11-bit weights)
    L.xtab_off = (int)xtab.size();
    L.ytab_off = (int)ytab.size();
    if (l > 0) {
      const LevelGeom& S = g.lv[l - 1];
      if (S.w == 2 * L.w && S.h == 2 * L.h) {
        set_error("exact 2x decimation: cv::resize switches INTER_LINEAR to INTER_AREA (not implemented)");
        return CUORB_E_UNSUPPORTED;
      }
      const double scale_x = 1. / ((double)L.w / S.w), scale_y = 1. / ((double)L.h / S.h);
      for (int dx = 0; dx < L.w; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = (int)std::floor(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= S.w - 1) { fx = 0; sx = S.w - 1; }
        xtab.push_back(ResizeTap{(short)sx, sat_short((1.f - fx) * 2048.f), sat_short(fx * 2048.f), 0});
      }
      for (int dy = 0; dy < L.h; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = (int)std::floor(fy);
        fy -= sy;
        ytab.push_back(ResizeTap{(short)sy, sat_short((1.f - fy) * 2048.f), sat_short(fy * 2048.f), 0});
      }
    }
  }
  g.total_cells     = cell_base;
  g.kp_cap          = key_off;
  g.key_img_stride  = key_off;
  g.pyr_img_stride  = pyr_off;
  g.blur_img_stride = blur_off;
  g.cand_img_stride = cand_off;
  if (g.kp_cap >= 65536) { set_error("more than 65535 keypoints per image"); return CUORB_E_UNSUPPORTED; }
  return CUORB_OK;
}

}  // namespace cuorb

// TMA tensor-map encoding through the driver entry point (no link-time dependency on libcuda)
namespace cuorb {
int tma_encode_u8_3d(CUtensorMap* out, void* base, uint64_t width_bytes, uint64_t rows, uint64_t images,
                     uint64_t row_pitch, uint64_t image_stride, uint32_t box_w, uint32_t box_h) {
  typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static encode_fn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) return -1;
    fn = (encode_fn)p;
  }
  const cuuint64_t dims[3] = {width_bytes, rows, images};
  const cuuint64_t strides[2] = {row_pitch, image_stride};
  const cuuint32_t box[3] = {box_w, box_h, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  return fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS
             ? 0 : -1;
}
}  // namespace cuorb

// ------------------------------------------------------------------------------------------
// handle lifetime
// ------------------------------------------------------------------------------------------
namespace cuorb {

template <typename T>
static cudaError_t dev_alloc(T** p, size_t count) {
  return cudaMalloc((void**)p, std::max<size_t>(count, 1) * sizeof(T));
}

static void free_all(cuorb_extractor* e) {
  cudaSetDevice(e->cfg.device);
  void* dev[] = {e->d_geom, e->d_xtab, e->d_ytab, e->d_wtab, e->d_blur_tiles, e->d_fast_cells, e->d_fast_recs, e->d_input, e->d_pyr, e->d_blur,
                 e->d_cand, e->d_cand_scratch, e->d_cand_count, e->d_key_count, e->d_pack, e->d_slot,
                 e->d_keys, e->d_sad, e->d_best_idx, e->d_best_dist, e->d_rtab, e->d_rstart};
  for (void* p : dev) if (p) cudaFree(p);
  void* host[] = {e->h_out_n, e->h_out_mono, e->h_flags, e->h_pack};
  for (void* p : host) if (p) cudaFreeHost(p);
  for (auto& ev : e->ev_stage) if (ev) cudaEventDestroy(ev);
  if (e->ev_done) cudaEventDestroy(e->ev_done);
  if (e->ev_pyr) cudaEventDestroy(e->ev_pyr);
  if (e->ev_blur) cudaEventDestroy(e->ev_blur);
  if (e->side_stream) cudaStreamDestroy(e->side_stream);
  if (e->stream) cudaStreamDestroy(e->stream);
}

// (re)build geometry for an image size and upload it; buffers were sized for the maximum
static int set_size(cuorb_extractor* e, int w, int h) {
  if (w == e->cur_w && h == e->cur_h) return CUORB_OK;
  if (w > e->cfg.max_width || h > e->cfg.max_height) {
    set_error("image larger than max_width x max_height given to cuorb_extractor_create");
    return CUORB_E_BADARG;
  }
  ExtractorGeom g;
  std::vector<ResizeTap> xt, yt;
  int rc = build_geometry(e, w, h, g, xt, yt);
  if (rc != CUORB_OK) return rc;
  std::vector<int4> tiles;
  build_blur_tiles(g, tiles);
  std::vector<int4> cells;
  build_fast_cells(g, cells);
  const size_t nb = (size_t)e->cfg.max_batch;
  if (g.pyr_img_stride * nb + 4096 > e->cap_pyr || g.blur_img_stride * nb > e->cap_blur ||
      g.cand_img_stride * nb > e->cap_cand || (size_t)g.key_img_stride * nb > e->cap_keys ||
      (size_t)g.kp_cap * nb > e->cap_kp || xt.size() > e->cap_xtab || yt.size() > e->cap_ytab ||
      tiles.size() > e->cap_tiles || cells.size() > e->cap_cells) {
    set_error("internal: geometry of this size exceeds the arena sized for max_width x max_height");
    return CUORB_E_CAPACITY;
  }
  // the previous geometry may still be in use by enqueued work
  CUORB_CUDA_TRY(cudaStreamSynchronize(e->stream));
  CUORB_CUDA_TRY(cudaMemcpy(e->d_geom, &g, sizeof(g), cudaMemcpyHostToDevice));
  if (!xt.empty()) CUORB_CUDA_TRY(cudaMemcpy(e->d_xtab, xt.data(), xt.size() * sizeof(ResizeTap), cudaMemcpyHostToDevice));
  if (!yt.empty()) CUORB_CUDA_TRY(cudaMemcpy(e->d_ytab, yt.data(), yt.size() * sizeof(ResizeTap), cudaMemcpyHostToDevice));
  {
    std::vector<uint4> wt;
    build_resize_words(g, xt, wt, e->wtab_off);
    if (wt.size() > e->cap_wtab) { set_error("internal: resize word tables exceed the arena"); return CUORB_E_CAPACITY; }
    if (!wt.empty()) CUORB_CUDA_TRY(cudaMemcpy(e->d_wtab, wt.data(), wt.size() * sizeof(uint4), cudaMemcpyHostToDevice));
    build_pyr_plan(g, wt, e->wtab_off, yt, e->pyr_plan);
  }
  CUORB_CUDA_TRY(cudaMemcpy(e->d_blur_tiles, tiles.data(), tiles.size() * sizeof(int4), cudaMemcpyHostToDevice));
  e->n_blur_tiles = (int)tiles.size();
  CUORB_CUDA_TRY(cudaMemcpy(e->d_fast_cells, cells.data(), cells.size() * sizeof(int4), cudaMemcpyHostToDevice));
  e->fast_persist = fast_persist_supported(g) && getenv("CUORB_FAST_LEGACY") == nullptr;
  if (e->fast_persist) {
    std::vector<int4> recs;
    e->n_fast_recs = build_fast_recs(g, recs);
    if (recs.size() > 2 * e->cap_cells) { set_error("internal: FAST work records exceed the arena"); return CUORB_E_CAPACITY; }
    if (!recs.empty()) CUORB_CUDA_TRY(cudaMemcpy(e->d_fast_recs, recs.data(), recs.size() * sizeof(int4), cudaMemcpyHostToDevice));
    memset(&e->fmaps, 0, sizeof(e->fmaps));
    for (int l = 0; l < g.nlevels; l++) {
      const LevelGeom& L = g.lv[l];
      if (tma_encode_u8_3d(&e->fmaps.lv[l], e->d_pyr + L.plane_off, (uint64_t)L.pitch, (uint64_t)L.h + 2 * EDGE_THRESHOLD, nb,
                           (uint64_t)L.pitch, g.pyr_img_stride, FT_PITCH, (uint32_t)(g.max_cell_h + 6)) != 0) {
        set_error("cuTensorMapEncodeTiled failed (TMA tensor map for the FAST tiles)");
        return CUORB_E_CUDA;
      }
    }
  }
  {
    TmaMaps tm;
    memset(&tm, 0, sizeof(tm));
    for (int l = 0; l < g.nlevels; l++) {
      const LevelGeom& L = g.lv[l];
      if (tma_encode_u8_3d(&tm.pyr_ic[l], e->d_pyr + L.plane_off, (uint64_t)L.pitch, (uint64_t)L.h + 2 * EDGE_THRESHOLD, nb,
                           (uint64_t)L.pitch, g.pyr_img_stride, IC_BOX_W, IC_BOX_H) != 0 ||
          tma_encode_u8_3d(&tm.blur_br[l], e->d_blur + L.blur_off, (uint64_t)L.bpitch, (uint64_t)L.h, nb, (uint64_t)L.bpitch,
                           g.blur_img_stride, BR_BOX_W, BR_BOX_H) != 0) {
        set_error("cuTensorMapEncodeTiled failed (TMA tensor map for the pyramid planes)");
        return CUORB_E_CUDA;
      }
    }
    e->tmaps = tm;
  }
  e->g = g; e->cur_w = w; e->cur_h = h;
  e->have_batch = false; e->have_stereo = false;
  return CUORB_OK;
}

}  // namespace cuorb

extern "C" {

const char* cuorb_last_error(void) { return t_last_error.c_str(); }
int cuorb_version(void) { return CUORB_VERSION; }
int cuorb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { set_error("cudaGetDeviceCount failed (no CUDA device / driver)"); return 0; }
  return n;
}
uint64_t cuorb_kernel_launch_count(void) { return g_launch_count.load(); }

void* cuorb_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaMallocHost(&p, bytes) != cudaSuccess) { set_error("cudaMallocHost failed"); return nullptr; }
  return p;
}
void cuorb_host_free(void* p) { if (p) cudaFreeHost(p); }

cuorb_extractor* cuorb_extractor_create(const cuorb_config* cfg) {
  if (!cfg || cfg->nlevels < 1 || cfg->nlevels > CUORB_MAX_LEVELS || cfg->nfeatures < 1 ||
      cfg->scale_factor <= 1.0f || cfg->max_width < 1 || cfg->max_height < 1 || cfg->max_batch < 1 ||
      cfg->max_batch > 65535) {
    set_error("cuorb_extractor_create: bad configuration");
    return nullptr;
  }
  if (cudaSetDevice(cfg->device) != cudaSuccess) {
    set_error("cuorb_extractor_create: cudaSetDevice failed (no usable CUDA device; there is no CPU fallback)");
    return nullptr;
  }
  cuorb_extractor* e = new cuorb_extractor();
  e->cfg = *cfg;
  build_tables(e);
  ExtractorGeom g;
  std::vector<ResizeTap> xt, yt;
  if (build_geometry(e, cfg->max_width, cfg->max_height, g, xt, yt) != CUORB_OK) { delete e; return nullptr; }
  std::vector<int4> tiles;
  build_blur_tiles(g, tiles);
  const size_t nb = (size_t)cfg->max_batch;
  e->cap_input = (size_t)cfg->max_width * cfg->max_height * nb;
  e->cap_pyr   = g.pyr_img_stride * nb + 4096;  // slack: tile loaders may read a few bytes past a plane
  e->cap_blur  = g.blur_img_stride * nb + 256;
  e->cap_cand  = g.cand_img_stride * nb;
  e->cap_keys  = (size_t)g.key_img_stride * nb;
  e->cap_kp    = (size_t)g.kp_cap * nb;
  e->cap_xtab  = xt.size() + 64; e->cap_ytab = yt.size() + 64; e->cap_tiles = tiles.size() + 64;
  e->cap_cells = (size_t)g.total_cells + 64;
  e->cap_wtab  = 64;
  for (int l = 1; l < g.nlevels; l++) e->cap_wtab += 2 * (size_t)(g.lv[l].pitch / 4 + 1);
  bool ok = true;
  ok &= cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) == cudaSuccess;
  ok &= cudaStreamCreateWithFlags(&e->side_stream, cudaStreamNonBlocking) == cudaSuccess;
  ok &= cudaEventCreateWithFlags(&e->ev_done, cudaEventDisableTiming) == cudaSuccess;
  ok &= cudaEventCreateWithFlags(&e->ev_pyr, cudaEventDisableTiming) == cudaSuccess;
  ok &= cudaEventCreateWithFlags(&e->ev_blur, cudaEventDisableTiming) == cudaSuccess;
  for (auto& ev : e->ev_stage) ok &= cudaEventCreate(&ev) == cudaSuccess;
  ok &= dev_alloc(&e->d_geom, 1) == cudaSuccess;
  ok &= dev_alloc(&e->d_xtab, e->cap_xtab) == cudaSuccess;
  ok &= dev_alloc(&e->d_wtab, e->cap_wtab) == cudaSuccess;
  ok &= dev_alloc(&e->d_ytab, e->cap_ytab) == cudaSuccess;
  ok &= dev_alloc(&e->d_blur_tiles, e->cap_tiles) == cudaSuccess;
  ok &= dev_alloc(&e->d_fast_cells, e->cap_cells) == cudaSuccess;
  ok &= dev_alloc(&e->d_fast_recs, 2 * e->cap_cells) == cudaSuccess;
  ok &= dev_alloc(&e->d_input, e->cap_input) == cudaSuccess;
  ok &= dev_alloc(&e->d_pyr, e->cap_pyr) == cudaSuccess;
  ok &= dev_alloc(&e->d_blur, e->cap_blur) == cudaSuccess;
  ok &= dev_alloc(&e->d_cand, e->cap_cand) == cudaSuccess;
  ok &= dev_alloc(&e->d_cand_scratch, 2 * e->cap_cand) == cudaSuccess;
  ok &= dev_alloc(&e->d_cand_count, nb * MAX_LEVELS) == cudaSuccess;
  ok &= dev_alloc(&e->d_key_count, nb * MAX_LEVELS) == cudaSuccess;
  ok &= dev_alloc(&e->d_slot, e->cap_kp) == cudaSuccess;
  ok &= dev_alloc(&e->d_keys, e->cap_keys) == cudaSuccess;
  {
    // result pack: [flags (2 ints) | n[nb] | mono[nb] | kps[cap_kp] | desc[cap_kp * 32] | uRight[cap_kp] | depth[cap_kp]],
    // sections 256-byte aligned
    e->pack_n_off    = 256;
    e->pack_mono_off = e->pack_n_off + align_up(nb * sizeof(int), 256);
    e->pack_kps_off  = e->pack_mono_off + align_up(nb * sizeof(int), 256);
    e->pack_desc_off = e->pack_kps_off + align_up(e->cap_kp * sizeof(cuorb_keypoint), 256);
    e->pack_ur_off   = e->pack_desc_off + align_up(e->cap_kp * 32, 256);
    e->pack_depth_off = e->pack_ur_off + align_up(e->cap_kp * sizeof(float), 256);
    e->pack_bytes    = e->pack_depth_off + align_up(e->cap_kp * sizeof(float), 256);
    ok &= dev_alloc(&e->d_pack, e->pack_bytes) == cudaSuccess;
    if (ok) {
      e->d_flags    = (int*)e->d_pack;
      e->d_out_n    = (int*)(e->d_pack + e->pack_n_off);
      e->d_out_mono = (int*)(e->d_pack + e->pack_mono_off);
      e->d_out_kps  = (cuorb_keypoint*)(e->d_pack + e->pack_kps_off);
      e->d_out_desc = e->d_pack + e->pack_desc_off;
      e->d_uright   = (float*)(e->d_pack + e->pack_ur_off);
      e->d_depth    = (float*)(e->d_pack + e->pack_depth_off);
    }
    if (nb == 1) ok &= cudaMallocHost((void**)&e->h_pack, e->pack_bytes) == cudaSuccess;
  }
  ok &= dev_alloc(&e->d_sad, e->cap_kp) == cudaSuccess;
  ok &= dev_alloc(&e->d_best_idx, e->cap_kp) == cudaSuccess;
  ok &= dev_alloc(&e->d_rtab, e->cap_kp) == cudaSuccess;
  ok &= dev_alloc(&e->d_rstart, nb * (STEREO_MAX_BUCKETS + 1)) == cudaSuccess;
  ok &= dev_alloc(&e->d_best_dist, e->cap_kp) == cudaSuccess;
  ok &= cudaMallocHost((void**)&e->h_out_n, nb * sizeof(int)) == cudaSuccess;
  ok &= cudaMallocHost((void**)&e->h_out_mono, nb * sizeof(int)) == cudaSuccess;
  ok &= cudaMallocHost((void**)&e->h_flags, sizeof(int)) == cudaSuccess;
  ok = ok && init_describe_constants() == 0;
  if (ok) ok = cudaMemset(e->d_pyr, 0, e->cap_pyr) == cudaSuccess;
  if (!ok || set_size(e, cfg->max_width, cfg->max_height) != CUORB_OK) {
    if (!ok) set_error(std::string("cuorb_extractor_create: CUDA allocation/initialisation failed: ") +
                       cudaGetErrorString(cudaGetLastError()));
    free_all(e);
    delete e;
    return nullptr;
  }
  return e;
}

void cuorb_extractor_destroy(cuorb_extractor* e) {
  if (!e) return;
  cudaSetDevice(e->cfg.device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  free_all(e);
  delete e;
}

int cuorb_extractor_get_tables(const cuorb_extractor* e, float* scale, float* inv_scale, float* sigma2,
                               float* inv_sigma2, int32_t* per_level) {
  if (!e) { set_error("null handle"); return CUORB_E_BADARG; }
  for (int i = 0; i < e->cfg.nlevels; i++) {
    if (scale) scale[i] = e->scale[i];
    if (inv_scale) inv_scale[i] = e->inv_scale[i];
    if (sigma2) sigma2[i] = e->sigma2[i];
    if (inv_sigma2) inv_sigma2[i] = e->inv_sigma2[i];
    if (per_level) per_level[i] = e->quota[i];
  }
  return e->cfg.nlevels;
}

int cuorb_extractor_capacity(const cuorb_extractor* e) {
  if (!e) { set_error("null handle"); return CUORB_E_BADARG; }
  return e->g.kp_cap;
}

void* cuorb_extractor_stream(cuorb_extractor* e) { return e ? (void*)e->stream : nullptr; }

int cuorb_set_profiling(cuorb_extractor* e, int enabled) {
  if (!e) { set_error("null handle"); return CUORB_E_BADARG; }
  e->profiling = enabled != 0;
  return CUORB_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// extraction
// ------------------------------------------------------------------------------------------
namespace cuorb {
static int check_flags(cuorb_extractor* e) {
  const int f = *e->h_flags;
  if (f & FLAG_CAND_OVERFLOW) { set_error("internal FAST candidate list overflow"); return CUORB_E_CAPACITY; }
  if (f & (FLAG_NODE_OVERFLOW | FLAG_KEY_OVERFLOW)) { set_error("internal quadtree node pool overflow"); return CUORB_E_CAPACITY; }
  return CUORB_OK;
}
}  // namespace cuorb

extern "C" {

// enqueue pyramid .. describe on device-resident gray planes (the H2D stage event is the caller's)
static int enqueue_pipeline(cuorb_extractor* e, int n_images, const uint8_t* dimgs, int width, int height,
                            size_t dstride, size_t dimg_stride, int lap0, int lap1) {
  cudaStream_t st = e->stream;
  const bool prof = e->profiling;
  // one or two images: the kernels are a few microseconds each, launch them as programmatic dependents (common.cuh)
  struct PdlScope { PdlScope(bool on) { ::cuorb::t_pdl = on; } ~PdlScope() { ::cuorb::t_pdl = false; } } pdl_scope(n_images <= 2 && !prof && pdl_allowed());
  // [0] overflow flags, [1] FAST work counter; per-level candidate counters.  One or two images: cleared by the
  // pyramid's first kernel (two memsets in front of the chain cost 8 us of a 130 us call)
  const bool clear_in_kernel = n_images <= 2;
  if (!clear_in_kernel) {
    CUORB_CUDA_TRY(cudaMemsetAsync(e->d_cand_count, 0, sizeof(int) * (size_t)n_images * e->g.nlevels, st));
    CUORB_CUDA_TRY(cudaMemsetAsync(e->d_flags, 0, 2 * sizeof(int), st));
  }
  if (prof) cudaEventRecord(e->ev_stage[CUORB_STAGE_PYRAMID], st);
  launch_pyramid(e->g, e->d_geom, dimgs, width, height, dstride, dimg_stride, n_images, e->d_pyr, e->d_xtab, e->d_ytab,
                 e->d_wtab, e->wtab_off, e->pyr_plan, st, clear_in_kernel ? e->d_cand_count : nullptr,
                 clear_in_kernel ? e->d_flags : nullptr);
  // FAST (integer-ALU bound, 20 warps per SM) and the blur (FMA pipe + loads) both depend only on the
  // pyramid: outside profiling runs the blur goes to the side stream, after FAST in launch order so
  // FAST's CTAs get their shared memory first and blur CTAs fill the remaining warp slots of each SM
  static const bool side_blur = []() { const char* v = getenv("CUORB_SIDE_BLUR"); return !v || atoi(v) != 0; }();
  const bool fork = side_blur && !prof;
  if (fork) CUORB_CUDA_TRY(cudaEventRecord(e->ev_pyr, st));
  if (prof) cudaEventRecord(e->ev_stage[CUORB_STAGE_FAST], st);
  if (e->fast_persist) {
    if (e->n_fast_recs > 0)
      launch_fast_persist(e->g, e->fmaps, e->d_fast_recs, e->n_fast_recs, n_images, e->d_cand, e->d_cand_count, e->d_flags,
                          e->d_flags + 1, st);
  } else {
    launch_fast(e->g, e->d_geom, e->d_fast_cells, n_images, e->d_pyr, e->d_cand, e->d_cand_count, e->d_flags, st);
  }
  if (fork) {
    CUORB_CUDA_TRY(cudaStreamWaitEvent(e->side_stream, e->ev_pyr, 0));
    launch_blur(&e->g, e->d_geom, n_images, e->d_blur_tiles, e->n_blur_tiles, e->d_pyr, e->d_blur, e->side_stream);
    CUORB_CUDA_TRY(cudaEventRecord(e->ev_blur, e->side_stream));
  }
  if (prof) cudaEventRecord(e->ev_stage[CUORB_STAGE_OCTREE], st);
  launch_octree(e->g, e->d_geom, n_images, e->d_cand, e->d_cand_scratch, e->d_cand_count, e->d_keys,
                e->d_key_count, e->d_flags, st);
  if (prof) cudaEventRecord(e->ev_stage[CUORB_STAGE_BLUR], st);
  if (fork) CUORB_CUDA_TRY(cudaStreamWaitEvent(st, e->ev_blur, 0));
  else launch_blur(&e->g, e->d_geom, n_images, e->d_blur_tiles, e->n_blur_tiles, e->d_pyr, e->d_blur, st);
  if (prof) cudaEventRecord(e->ev_stage[CUORB_STAGE_DESCRIBE], st);
  launch_describe(e->g, e->d_geom, e->tmaps, n_images, e->d_pyr, e->d_blur, e->d_keys, e->d_key_count, lap0, lap1,
                  e->d_slot, e->d_out_kps, e->d_out_desc, e->d_out_n, e->d_out_mono, st);
  if (prof) cudaEventRecord(e->ev_stage[CUORB_STAGE_STEREO], st);
  CUORB_CUDA_TRY(cudaGetLastError());
  CUORB_CUDA_TRY(cudaEventRecord(e->ev_done, st));
  e->n_img = n_images; e->have_batch = true; e->have_stereo = false;
  e->have_stage_times = prof; e->stereo_timed = false;
  return CUORB_OK;
}

int cuorb_extract_batch_async(cuorb_extractor* e, int n_images, const uint8_t* imgs, int width,
                              int height, size_t stride, size_t image_stride, cuorb_mem where,
                              int lap0, int lap1) {
  if (!e) { set_error("null handle"); return CUORB_E_BADARG; }
  if (!imgs || width <= 0 || height <= 0 || n_images <= 0) { set_error("empty image"); return CUORB_E_EMPTY; }
  if (n_images > e->cfg.max_batch || stride < (size_t)width) { set_error("batch larger than max_batch or stride < width"); return CUORB_E_BADARG; }
  CUORB_CUDA_TRY(cudaSetDevice(e->cfg.device));
  int rc = set_size(e, width, height);
  if (rc != CUORB_OK) return rc;
  cudaStream_t st = e->stream;
  if (e->profiling) cudaEventRecord(e->ev_stage[CUORB_STAGE_H2D], st);
  const uint8_t* dimgs = imgs;
  size_t dstride = stride, dimg_stride = image_stride;
  if (where == CUORB_MEM_HOST) {
    if (stride == (size_t)width && (n_images == 1 || image_stride == (size_t)width * height)) {
      CUORB_CUDA_TRY(cudaMemcpyAsync(e->d_input, imgs, (size_t)width * height * n_images, cudaMemcpyHostToDevice, st));
    } else {
      for (int i = 0; i < n_images; i++)
        CUORB_CUDA_TRY(cudaMemcpy2DAsync(e->d_input + (size_t)i * width * height, width, imgs + (size_t)i * image_stride,
                                         stride, width, height, cudaMemcpyHostToDevice, st));
    }
    dimgs = e->d_input; dstride = width; dimg_stride = (size_t)width * height;
  }
  return enqueue_pipeline(e, n_images, dimgs, width, height, dstride, dimg_stride, lap0, lap1);
}

// ------------------------------------------------------------------------------------------
// pre-extraction image operations
// ------------------------------------------------------------------------------------------
struct cuorb_preproc {
  int device = 0, mode = 0, sw = 0, sh = 0, dw = 0, dh = 0, max_batch = 1, area2 = 0;
  float *d_mapx = nullptr, *d_mapy = nullptr;
  int *  d_xofs = nullptr, *d_yofs = nullptr;
  short *d_xa = nullptr, *d_yb = nullptr;
  uint8_t *d_raw = nullptr, *d_gray = nullptr;  // staging: raw frames from the host / gray planes for the host
  cudaStream_t stream = nullptr;
};

static int pix_channels(int fmt) { return fmt == CUORB_PIX_GRAY8 ? 1 : (fmt <= CUORB_PIX_BGR8 ? 3 : 4); }

void cuorb_preproc_destroy(cuorb_preproc* p) {
  if (!p) return;
  cudaSetDevice(p->device);
  void* dev[] = {p->d_mapx, p->d_mapy, p->d_xofs, p->d_yofs, p->d_xa, p->d_yb, p->d_raw, p->d_gray};
  for (void* d : dev) if (d) cudaFree(d);
  if (p->stream) cudaStreamDestroy(p->stream);
  delete p;
}

cuorb_preproc* cuorb_preproc_create(int device, cuorb_pre_mode mode, int src_w, int src_h, int dst_w, int dst_h,
                                    const float* map_x, const float* map_y, int max_batch) {
  if (src_w <= 0 || src_h <= 0 || max_batch <= 0 || mode < CUORB_PRE_NONE || mode > CUORB_PRE_RESIZE) {
    set_error("cuorb_preproc_create: bad size / mode"); return nullptr;
  }
  if (mode == CUORB_PRE_NONE) { dst_w = src_w; dst_h = src_h; }
  if (dst_w <= 0 || dst_h <= 0 || (mode == CUORB_PRE_REMAP && (!map_x || !map_y))) {
    set_error("cuorb_preproc_create: bad destination size or missing maps"); return nullptr;
  }
  if (cudaSetDevice(device) != cudaSuccess) { set_error("cuorb_preproc_create: cudaSetDevice failed (no CUDA device; libcuorb has no CPU fallback)"); return nullptr; }
  cuorb_preproc* p = new cuorb_preproc();
  p->device = device; p->mode = mode; p->sw = src_w; p->sh = src_h; p->dw = dst_w; p->dh = dst_h; p->max_batch = max_batch;
  bool ok = cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking) == cudaSuccess;
  ok &= dev_alloc(&p->d_raw, (size_t)src_w * src_h * 4 * max_batch) == cudaSuccess;
  ok &= dev_alloc(&p->d_gray, (size_t)dst_w * dst_h * max_batch) == cudaSuccess;
  if (ok && mode == CUORB_PRE_REMAP) {
    const size_t n = (size_t)dst_w * dst_h;
    ok &= dev_alloc(&p->d_mapx, n) == cudaSuccess && dev_alloc(&p->d_mapy, n) == cudaSuccess;
    ok = ok && cudaMemcpy(p->d_mapx, map_x, n * sizeof(float), cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(p->d_mapy, map_y, n * sizeof(float), cudaMemcpyHostToDevice) == cudaSuccess;
  }
  if (ok && mode == CUORB_PRE_RESIZE) {
    p->area2 = (src_w == 2 * dst_w && src_h == 2 * dst_h) ? 1 : 0;
    // cv::resize INTER_LINEAR coefficient tables (imgproc/resize.cpp; same float sequence as the pyramid's)
    std::vector<int> xofs(dst_w), yofs(dst_h);
    std::vector<short> xa(2 * (size_t)dst_w), yb(2 * (size_t)dst_h);
    const double scale_x = 1. / ((double)dst_w / src_w), scale_y = 1. / ((double)dst_h / src_h);
    for (int dx = 0; dx < dst_w; dx++) {
      float fx = (float)((dx + 0.5) * scale_x - 0.5);
      int sx = (int)std::floor(fx);
      fx -= sx;
      if (sx < 0) { fx = 0; sx = 0; }
      if (sx >= src_w - 1) { fx = 0; sx = src_w - 1; }
      xofs[dx] = sx; xa[2 * dx] = sat_short((1.f - fx) * 2048.f); xa[2 * dx + 1] = sat_short(fx * 2048.f);
    }
    for (int dy = 0; dy < dst_h; dy++) {
      float fy = (float)((dy + 0.5) * scale_y - 0.5);
      int sy = (int)std::floor(fy);
      fy -= sy;
      yofs[dy] = sy; yb[2 * dy] = sat_short((1.f - fy) * 2048.f); yb[2 * dy + 1] = sat_short(fy * 2048.f);
    }
    ok &= dev_alloc(&p->d_xofs, xofs.size()) == cudaSuccess && dev_alloc(&p->d_yofs, yofs.size()) == cudaSuccess;
    ok &= dev_alloc(&p->d_xa, xa.size()) == cudaSuccess && dev_alloc(&p->d_yb, yb.size()) == cudaSuccess;
    ok = ok && cudaMemcpy(p->d_xofs, xofs.data(), xofs.size() * sizeof(int), cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(p->d_yofs, yofs.data(), yofs.size() * sizeof(int), cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(p->d_xa, xa.data(), xa.size() * sizeof(short), cudaMemcpyHostToDevice) == cudaSuccess &&
         cudaMemcpy(p->d_yb, yb.data(), yb.size() * sizeof(short), cudaMemcpyHostToDevice) == cudaSuccess;
  }
  if (!ok) {
    set_error(std::string("cuorb_preproc_create: ") + cudaGetErrorString(cudaGetLastError()));
    cuorb_preproc_destroy(p);
    return nullptr;
  }
  return p;
}

// src (host or device) -> gray planes at dst (device), on stream st
static int enqueue_preproc(cuorb_preproc* p, int n, const uint8_t* src, size_t stride, size_t image_stride,
                           cuorb_mem where, int fmt, uint8_t* d_dst, cudaStream_t st) {
  const int ch = pix_channels(fmt);
  const size_t row = (size_t)p->sw * ch;
  if (stride < row) { set_error("preprocess: stride smaller than one row"); return CUORB_E_BADARG; }
  const uint8_t* dsrc = src;
  size_t dstride = stride, dimg = image_stride;
  if (where == CUORB_MEM_HOST) {
    if (stride == row && (n == 1 || image_stride == row * p->sh)) {
      CUORB_CUDA_TRY(cudaMemcpyAsync(p->d_raw, src, row * p->sh * n, cudaMemcpyHostToDevice, st));
    } else {
      for (int i = 0; i < n; i++)
        CUORB_CUDA_TRY(cudaMemcpy2DAsync(p->d_raw + (size_t)i * row * p->sh, row, src + (size_t)i * image_stride, stride,
                                         row, p->sh, cudaMemcpyHostToDevice, st));
    }
    dsrc = p->d_raw; dstride = row; dimg = row * p->sh;
  }
  if (ch == 1 && p->mode == CUORB_PRE_NONE) {  // nothing to compute: the reference clones the frame
    for (int i = 0; i < n; i++)
      CUORB_CUDA_TRY(cudaMemcpy2DAsync(d_dst + (size_t)i * p->dw * p->dh, p->dw, dsrc + (size_t)i * dimg, dstride, p->dw,
                                       p->dh, cudaMemcpyDeviceToDevice, st));
    return CUORB_OK;
  }
  PreArgs a;
  a.src = dsrc; a.sstride = dstride; a.simg_stride = dimg; a.sw = p->sw; a.sh = p->sh;
  a.dst = d_dst; a.dstride = p->dw; a.dimg_stride = (size_t)p->dw * p->dh; a.dw = p->dw; a.dh = p->dh;
  a.mapx = p->d_mapx; a.mapy = p->d_mapy; a.xofs = p->d_xofs; a.xa = p->d_xa; a.yofs = p->d_yofs; a.yb = p->d_yb;
  a.area2 = p->area2;
  const bool rgb = fmt == CUORB_PIX_RGB8 || fmt == CUORB_PIX_RGBA8;
  a.ir = rgb ? 0 : 2; a.ib = rgb ? 2 : 0;
  launch_preproc(a, ch, p->mode, n, st);
  CUORB_CUDA_TRY(cudaGetLastError());
  return CUORB_OK;
}

static int preproc_args_ok(cuorb_preproc* p, int n, const uint8_t* src, int fmt) {
  if (!p) { set_error("null preproc handle"); return CUORB_E_BADARG; }
  if (!src || n <= 0) { set_error("empty image"); return CUORB_E_EMPTY; }
  if (n > p->max_batch || fmt < CUORB_PIX_GRAY8 || fmt > CUORB_PIX_BGRA8) { set_error("batch larger than max_batch or unknown pixel format"); return CUORB_E_BADARG; }
  return CUORB_OK;
}

int cuorb_preprocess(cuorb_preproc* p, int n, const uint8_t* src, size_t stride, size_t image_stride,
                     cuorb_mem src_where, cuorb_pixfmt fmt, uint8_t* dst, cuorb_mem dst_where) {
  int rc = preproc_args_ok(p, n, src, fmt);
  if (rc != CUORB_OK) return rc;
  if (!dst) { set_error("null destination"); return CUORB_E_BADARG; }
  CUORB_CUDA_TRY(cudaSetDevice(p->device));
  uint8_t* d_dst = dst_where == CUORB_MEM_DEVICE ? dst : p->d_gray;
  rc = enqueue_preproc(p, n, src, stride, image_stride, src_where, fmt, d_dst, p->stream);
  if (rc != CUORB_OK) return rc;
  if (dst_where == CUORB_MEM_HOST)
    CUORB_CUDA_TRY(cudaMemcpyAsync(dst, p->d_gray, (size_t)p->dw * p->dh * n, cudaMemcpyDeviceToHost, p->stream));
  CUORB_CUDA_TRY(cudaStreamSynchronize(p->stream));
  return n;
}

int cuorb_extract_batch_async_pre(cuorb_extractor* e, cuorb_preproc* p, int n, const uint8_t* src, size_t stride,
                                  size_t image_stride, cuorb_mem where, cuorb_pixfmt fmt, int lap0, int lap1) {
  if (!e) { set_error("null handle"); return CUORB_E_BADARG; }
  int rc = preproc_args_ok(p, n, src, fmt);
  if (rc != CUORB_OK) return rc;
  if (n > e->cfg.max_batch || p->device != e->cfg.device) { set_error("batch larger than the extractor's max_batch, or handles on different devices"); return CUORB_E_BADARG; }
  CUORB_CUDA_TRY(cudaSetDevice(e->cfg.device));
  rc = set_size(e, p->dw, p->dh);
  if (rc != CUORB_OK) return rc;
  if (e->profiling) cudaEventRecord(e->ev_stage[CUORB_STAGE_H2D], e->stream);
  rc = enqueue_preproc(p, n, src, stride, image_stride, where, fmt, e->d_input, e->stream);
  if (rc != CUORB_OK) return rc;
  return enqueue_pipeline(e, n, e->d_input, p->dw, p->dh, (size_t)p->dw, (size_t)p->dw * p->dh, lap0, lap1);
}

int cuorb_extract_fetch_async(cuorb_extractor* e, cuorb_keypoint* kps, uint8_t* desc, int capacity,
                              int32_t* n_out, int32_t* mono_out) {
  if (!e) { set_error("null handle"); return CUORB_E_BADARG; }
  if (!e->have_batch) { set_error("fetch before extract"); return CUORB_E_STATE; }
  if ((kps || desc) && capacity <= 0) { set_error("capacity <= 0"); return CUORB_E_BADARG; }
  CUORB_CUDA_TRY(cudaSetDevice(e->cfg.device));
  cudaStream_t st = e->stream;
  const int n = e->n_img, cap = e->g.kp_cap, cw = std::min(capacity, cap);
  e->fetch_n = n; e->fetch_capacity = (kps || desc) ? capacity : 0;
  CUORB_CUDA_TRY(cudaMemcpyAsync(e->h_out_n, e->d_out_n, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
  CUORB_CUDA_TRY(cudaMemcpyAsync(e->h_out_mono, e->d_out_mono, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
  CUORB_CUDA_TRY(cudaMemcpyAsync(e->h_flags, e->d_flags, sizeof(int), cudaMemcpyDeviceToHost, st));
  if (n_out) CUORB_CUDA_TRY(cudaMemcpyAsync(n_out, e->d_out_n, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
  if (mono_out) CUORB_CUDA_TRY(cudaMemcpyAsync(mono_out, e->d_out_mono, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
  if (kps)
    CUORB_CUDA_TRY(cudaMemcpy2DAsync(kps, (size_t)capacity * sizeof(cuorb_keypoint), e->d_out_kps,
                                     (size_t)cap * sizeof(cuorb_keypoint), (size_t)cw * sizeof(cuorb_keypoint), n,
                                     cudaMemcpyDeviceToHost, st));
  if (desc)
    CUORB_CUDA_TRY(cudaMemcpy2DAsync(desc, (size_t)capacity * 32, e->d_out_desc, (size_t)cap * 32, (size_t)cw * 32, n,
                                     cudaMemcpyDeviceToHost, st));
  return n;
}

int cuorb_sync(cuorb_extractor* e) {
  if (!e) { set_error("null handle"); return CUORB_E_BADARG; }
  CUORB_CUDA_TRY(cudaSetDevice(e->cfg.device));
  CUORB_CUDA_TRY(cudaStreamSynchronize(e->stream));
  return CUORB_OK;
}

// after a fetch has completed (stream synchronised): overflow flags and caller capacity
int cuorb_extractor_status(cuorb_extractor* e) {
  if (!e) { set_error("null handle"); return CUORB_E_BADARG; }
  if (e->fetch_n <= 0) return CUORB_OK;
  int rc = check_flags(e);
  if (rc != CUORB_OK) return rc;
  if (e->fetch_capacity > 0)
    for (int i = 0; i < e->fetch_n; i++)
      if (e->h_out_n[i] > e->fetch_capacity) { set_error("caller capacity smaller than the keypoint count"); return CUORB_E_CAPACITY; }
  return CUORB_OK;
}

int cuorb_extract_fetch(cuorb_extractor* e, cuorb_keypoint* kps, uint8_t* desc, int capacity,
                        int32_t* n_out, int32_t* mono_out) {
  int n = cuorb_extract_fetch_async(e, kps, desc, capacity, n_out, mono_out);
  if (n < 0) return n;
  int rc = cuorb_sync(e);
  if (rc != CUORB_OK) return rc;
  rc = cuorb_extractor_status(e);
  if (rc != CUORB_OK) return rc;
  return n;
}

int cuorb_extract(cuorb_extractor* e, const uint8_t* img, int width, int height, size_t stride,
                  int lap0, int lap1, cuorb_keypoint* kps, uint8_t* desc, int capacity, int* n_out) {
  int rc = cuorb_extract_batch_async(e, 1, img, width, height, stride, stride * (size_t)height, CUORB_MEM_HOST, lap0, lap1);
  if (rc != CUORB_OK) return rc;
  if (e->h_pack) {
    // one-image handle: flags, count, monoIndex, keypoints and descriptors come back in ONE copy into the pinned
    // mirror (62 KB at 1000 features) and are handed to the caller's buffers from there, instead of seven separate
    // copies, four of them staged through pageable memory
    const int cap = e->g.kp_cap;
    const size_t bytes = e->pack_desc_off + (size_t)cap * 32;
    CUORB_CUDA_TRY(cudaMemcpyAsync(e->h_pack, e->d_pack, bytes, cudaMemcpyDeviceToHost, e->stream));
    CUORB_CUDA_TRY(cudaStreamSynchronize(e->stream));
    const int n1 = *(const int*)(e->h_pack + e->pack_n_off), mono1 = *(const int*)(e->h_pack + e->pack_mono_off);
    *e->h_flags = *(const int*)e->h_pack;
    e->h_out_n[0] = n1; e->h_out_mono[0] = mono1;
    e->fetch_n = 1; e->fetch_capacity = (kps || desc) ? capacity : 0;
    if (n_out) *n_out = n1;
    rc = check_flags(e);
    if (rc != CUORB_OK) return rc;
    if ((kps || desc) && n1 > capacity) { set_error("caller capacity smaller than the keypoint count"); return CUORB_E_CAPACITY; }
    if (kps && n1 > 0) memcpy(kps, e->h_pack + e->pack_kps_off, sizeof(cuorb_keypoint) * (size_t)n1);
    if (desc && n1 > 0) memcpy(desc, e->h_pack + e->pack_desc_off, 32 * (size_t)n1);
    return mono1;
  }
  int32_t n = 0, mono = 0;
  rc = cuorb_extract_fetch(e, kps, desc, capacity, &n, &mono);
  if (n_out) *n_out = n;
  if (rc < 0) return rc;
  return mono;
}

int cuorb_extractor_pyramid(const cuorb_extractor* e, int image_index, int level, const uint8_t** dev_ptr,
                            int* width, int* height, size_t* stride) {
  if (!e || !e->have_batch || image_index < 0 || image_index >= e->n_img || level < 0 || level >= e->g.nlevels) {
    set_error("cuorb_extractor_pyramid: bad handle / index, or no batch extracted yet");
    return CUORB_E_BADARG;
  }
  const LevelGeom& L = e->g.lv[level];
  if (dev_ptr) *dev_ptr = e->d_pyr + (size_t)image_index * e->g.pyr_img_stride + L.pix_off;
  if (width) *width = L.w;
  if (height) *height = L.h;
  if (stride) *stride = (size_t)L.pitch;
  return CUORB_OK;
}

int cuorb_extractor_device_features(const cuorb_extractor* e, int image_index, const cuorb_keypoint** d_keys,
                                    const uint8_t** d_descriptors, const float** d_u_right, const float** d_depth) {
  if (!e || !e->have_batch || image_index < 0 || image_index >= e->n_img) {
    set_error("cuorb_extractor_device_features: bad handle / index, or no batch extracted yet");
    return CUORB_E_BADARG;
  }
  const size_t off = (size_t)image_index * e->g.kp_cap;
  if (d_keys) *d_keys = e->d_out_kps + off;
  if (d_descriptors) *d_descriptors = e->d_out_desc + off * 32;
  if (d_u_right) *d_u_right = e->have_stereo ? e->d_uright + off : nullptr;
  if (d_depth) *d_depth = e->have_stereo ? e->d_depth + off : nullptr;
  return e->g.kp_cap;
}

int64_t cuorb_extractor_download_plane(cuorb_extractor* e, int image_index, int level, int plane,
                                       uint8_t* out, size_t out_bytes, int* width, int* height) {
  if (!e || !e->have_batch || image_index < 0 || image_index >= e->n_img || level < 0 ||
      level >= e->g.nlevels || plane < 0 || plane > 2 || !out) {
    set_error("cuorb_extractor_download_plane: bad argument");
    return CUORB_E_BADARG;
  }
  CUORB_CUDA_TRY(cudaSetDevice(e->cfg.device));
  const LevelGeom& L = e->g.lv[level];
  if (width) *width = L.w;
  if (height) *height = L.h;
  const int B = EDGE_THRESHOLD;
  const size_t ow = plane == 0 ? L.w + 2 * B : L.w, oh = plane == 0 ? L.h + 2 * B : L.h;
  if (out_bytes < ow * oh) { set_error("output buffer too small"); return CUORB_E_CAPACITY; }
  CUORB_CUDA_TRY(cudaStreamSynchronize(e->stream));
  const uint8_t* src; size_t spitch;
  if (plane == 0) { src = e->d_pyr + (size_t)image_index * e->g.pyr_img_stride + L.plane_off + (ROW_LEAD - B); spitch = L.pitch; }
  else if (plane == 1) { src = e->d_pyr + (size_t)image_index * e->g.pyr_img_stride + L.pix_off; spitch = L.pitch; }
  else { src = e->d_blur + (size_t)image_index * e->g.blur_img_stride + L.blur_off; spitch = L.bpitch; }
  CUORB_CUDA_TRY(cudaMemcpy2D(out, ow, src, spitch, ow, oh, cudaMemcpyDeviceToHost));
  return (int64_t)(ow * oh);
}

int cuorb_extractor_download_candidates(cuorb_extractor* e, int image_index, int level, int32_t* xs,
                                        int32_t* ys, int32_t* responses, int capacity) {
  if (!e || !e->have_batch || image_index < 0 || image_index >= e->n_img || level < 0 || level >= e->g.nlevels) {
    set_error("cuorb_extractor_download_candidates: bad argument");
    return CUORB_E_BADARG;
  }
  // the device list is unordered (appended with atomics); sorted here into the reference's order
  CUORB_CUDA_TRY(cudaSetDevice(e->cfg.device));
  CUORB_CUDA_TRY(cudaStreamSynchronize(e->stream));
  const LevelGeom& L = e->g.lv[level];
  int cnt = 0;
  CUORB_CUDA_TRY(cudaMemcpy(&cnt, e->d_cand_count + image_index * e->g.nlevels + level, sizeof(int), cudaMemcpyDeviceToHost));
  cnt = std::min(cnt, L.cand_cap);
  std::vector<unsigned> a(cnt);
  if (cnt) {
    CUORB_CUDA_TRY(cudaMemcpy(a.data(), e->d_cand + (size_t)image_index * e->g.cand_img_stride + L.cand_off,
                              sizeof(unsigned) * cnt, cudaMemcpyDeviceToHost));
  }
  std::vector<unsigned long long> ord(cnt);
  for (int i = 0; i < cnt; i++) {
    const int x = cand_x(a[i]), y = cand_y(a[i]);
    const unsigned long long cell = (unsigned long long)((y - 3) / L.hCell) * L.nCols + (x - 3) / L.wCell;
    ord[i] = (cell << 40) | ((unsigned long long)y << 28) | ((unsigned long long)x << 16) | (unsigned long long)cand_s(a[i]);
  }
  std::sort(ord.begin(), ord.end());
  for (int i = 0; i < cnt && i < capacity; i++) {
    xs[i] = (int)((ord[i] >> 16) & 0xFFF); ys[i] = (int)((ord[i] >> 28) & 0xFFF); responses[i] = (int)(ord[i] & 0xFF);
  }
  return cnt;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// stereo association
// ------------------------------------------------------------------------------------------
extern "C" {

int cuorb_stereo_match_batch_async(cuorb_extractor* l, cuorb_extractor* r, float mb, float mbf) {
  if (!l || !r) { set_error("null handle"); return CUORB_E_BADARG; }
  if (!l->have_batch || !r->have_batch || l->n_img != r->n_img) { set_error("stereo match needs one extracted batch of equal size on both handles"); return CUORB_E_STATE; }
  if (l->cfg.device != r->cfg.device || l->cur_w != r->cur_w || l->cur_h != r->cur_h ||
      l->g.nlevels != r->g.nlevels || l->g.kp_cap != r->g.kp_cap || l->cfg.scale_factor != r->cfg.scale_factor) {
    set_error("stereo match needs two extractors with identical configuration on one device");
    return CUORB_E_BADARG;
  }
  CUORB_CUDA_TRY(cudaSetDevice(l->cfg.device));
  CUORB_CUDA_TRY(cudaStreamWaitEvent(l->stream, r->ev_done, 0));
  StereoArgs a;
  a.dgL = l->d_geom; a.pyrL = l->d_pyr; a.pyrR = r->d_pyr;
  a.kpsL = l->d_out_kps; a.kpsR = r->d_out_kps; a.descL = l->d_out_desc; a.descR = r->d_out_desc;
  a.nL = l->d_out_n; a.nR = r->d_out_n; a.kp_cap = l->g.kp_cap; a.n_img = l->n_img;
  a.mb = mb; a.mbf = mbf; a.u_right = l->d_uright; a.depth = l->d_depth;
  a.best_idx_r = l->d_best_idx; a.best_dist = l->d_best_dist; a.sad = l->d_sad;
  a.rtab = l->d_rtab; a.rstart = l->d_rstart;
  launch_stereo(l->g, a, l->stream);
  if (l->profiling) { cudaEventRecord(l->ev_stage[CUORB_STAGE_COUNT], l->stream); l->stereo_timed = true; }
  CUORB_CUDA_TRY(cudaGetLastError());
  CUORB_CUDA_TRY(cudaEventRecord(l->ev_done, l->stream));
  // the match reads the right handle's pyramid, keypoints and descriptors: whatever is enqueued on the
  // right handle next (its next extraction overwrites them) has to wait for it
  CUORB_CUDA_TRY(cudaStreamWaitEvent(r->stream, l->ev_done, 0));
  l->have_stereo = true;
  return CUORB_OK;
}

int cuorb_stereo_fetch_async(cuorb_extractor* l, float* u_right, float* depth, int capacity) {
  if (!l) { set_error("null handle"); return CUORB_E_BADARG; }
  if (!l->have_stereo) { set_error("stereo fetch before stereo match"); return CUORB_E_STATE; }
  CUORB_CUDA_TRY(cudaSetDevice(l->cfg.device));
  const int n = l->n_img, cap = l->g.kp_cap, cw = std::min(capacity, cap);
  if (u_right) CUORB_CUDA_TRY(cudaMemcpy2DAsync(u_right, (size_t)capacity * 4, l->d_uright, (size_t)cap * 4, (size_t)cw * 4, n, cudaMemcpyDeviceToHost, l->stream));
  if (depth) CUORB_CUDA_TRY(cudaMemcpy2DAsync(depth, (size_t)capacity * 4, l->d_depth, (size_t)cap * 4, (size_t)cw * 4, n, cudaMemcpyDeviceToHost, l->stream));
  return n;
}

int cuorb_stereo_fetch(cuorb_extractor* l, float* u_right, float* depth, int capacity,
                       int32_t* best_idx_r, int32_t* best_dist, int32_t* sad) {
  int n = cuorb_stereo_fetch_async(l, u_right, depth, capacity);
  if (n < 0) return n;
  const int cap = l->g.kp_cap, cw = std::min(capacity, cap);
  int32_t* dst[3] = {best_idx_r, best_dist, sad};
  int* src[3] = {l->d_best_idx, l->d_best_dist, l->d_sad};
  for (int k = 0; k < 3; k++)
    if (dst[k]) CUORB_CUDA_TRY(cudaMemcpy2DAsync(dst[k], (size_t)capacity * 4, src[k], (size_t)cap * 4, (size_t)cw * 4, n, cudaMemcpyDeviceToHost, l->stream));
  int rc = cuorb_sync(l);
  return rc != CUORB_OK ? rc : n;
}

// Host-pointer form: uploads the caller's keypoints/descriptors into the handles' result slots
// of image 0 (they normally already are there) and runs the same kernels.
int cuorb_stereo_match(cuorb_extractor* l, cuorb_extractor* r, int n_left, const cuorb_keypoint* kps_left,
                       const uint8_t* desc_left, int n_right, const cuorb_keypoint* kps_right,
                       const uint8_t* desc_right, float mb, float mbf, float* u_right, float* depth) {
  if (!l || !r || !u_right || !depth) { set_error("null argument"); return CUORB_E_BADARG; }
  if (!l->have_batch || !r->have_batch) { set_error("stereo match needs the pyramids of a previous extract on both handles"); return CUORB_E_STATE; }
  if (n_left < 0 || n_right < 0 || n_left > l->g.kp_cap || n_right > r->g.kp_cap) { set_error("keypoint count beyond handle capacity"); return CUORB_E_CAPACITY; }
  CUORB_CUDA_TRY(cudaSetDevice(l->cfg.device));
  if (n_left) {
    CUORB_CUDA_TRY(cudaMemcpyAsync(l->d_out_kps, kps_left, sizeof(cuorb_keypoint) * n_left, cudaMemcpyHostToDevice, l->stream));
    CUORB_CUDA_TRY(cudaMemcpyAsync(l->d_out_desc, desc_left, 32 * (size_t)n_left, cudaMemcpyHostToDevice, l->stream));
  }
  CUORB_CUDA_TRY(cudaMemcpyAsync(l->d_out_n, &n_left, sizeof(int), cudaMemcpyHostToDevice, l->stream));
  if (n_right) {
    CUORB_CUDA_TRY(cudaMemcpyAsync(r->d_out_kps, kps_right, sizeof(cuorb_keypoint) * n_right, cudaMemcpyHostToDevice, r->stream));
    CUORB_CUDA_TRY(cudaMemcpyAsync(r->d_out_desc, desc_right, 32 * (size_t)n_right, cudaMemcpyHostToDevice, r->stream));
  }
  CUORB_CUDA_TRY(cudaMemcpyAsync(r->d_out_n, &n_right, sizeof(int), cudaMemcpyHostToDevice, r->stream));
  CUORB_CUDA_TRY(cudaStreamSynchronize(l->stream));  // &n_left / &n_right are stack variables
  CUORB_CUDA_TRY(cudaStreamSynchronize(r->stream));
  CUORB_CUDA_TRY(cudaEventRecord(r->ev_done, r->stream));
  struct Restore {  // the batch size of both handles is put back on every exit path
    cuorb_extractor *l, *r; int nl, nr;
    ~Restore() { l->n_img = nl; r->n_img = nr; }
  } restore{l, r, l->n_img, r->n_img};
  l->n_img = 1; r->n_img = 1;
  int rc = cuorb_stereo_match_batch_async(l, r, mb, mbf);
  if (rc == CUORB_OK) {
    CUORB_CUDA_TRY(cudaMemcpyAsync(u_right, l->d_uright, 4 * (size_t)n_left, cudaMemcpyDeviceToHost, l->stream));
    CUORB_CUDA_TRY(cudaMemcpyAsync(depth, l->d_depth, 4 * (size_t)n_left, cudaMemcpyDeviceToHost, l->stream));
    rc = cuorb_sync(l);
  }
  return rc;
}

// Fused stereo frame (SURVEY 8b "fused cuorb_extract_stereo"): what the stereo Frame constructor does in three
// steps - two ExtractORB threads (src/Frame.cc:202-205, lapping {0,0}) and ComputeStereoMatches (:231) - as one call.
// Both extractions are enqueued on their handles' own streams, the association runs on the device on the resident
// keypoints / descriptors / pyramids (nothing is re-uploaded), and each handle's results come back in one copy.
int cuorb_extract_stereo(cuorb_extractor* l, cuorb_extractor* r, const uint8_t* img_left, const uint8_t* img_right,
                         int width, int height, size_t stride, float mb, float mbf, cuorb_keypoint* kps_left,
                         uint8_t* desc_left, int* n_left, cuorb_keypoint* kps_right, uint8_t* desc_right, int* n_right,
                         int capacity, float* u_right, float* depth) {
  if (!l || !r || l == r) { set_error("cuorb_extract_stereo: two distinct handles needed"); return CUORB_E_BADARG; }
  if (!l->h_pack || !r->h_pack) { set_error("cuorb_extract_stereo: both handles must be created with max_batch == 1"); return CUORB_E_BADARG; }
  if (!kps_left || !desc_left || !kps_right || !desc_right || !n_left || !n_right || !u_right || !depth || capacity <= 0) {
    set_error("cuorb_extract_stereo: null output"); return CUORB_E_BADARG;
  }
  int rc = cuorb_extract_batch_async(l, 1, img_left, width, height, stride, stride * (size_t)height, CUORB_MEM_HOST, 0, 0);
  if (rc != CUORB_OK) return rc;
  rc = cuorb_extract_batch_async(r, 1, img_right, width, height, stride, stride * (size_t)height, CUORB_MEM_HOST, 0, 0);
  if (rc != CUORB_OK) return rc;
  // the right handle's results are complete at its ev_done: fetch them on its own stream, beside the match
  const size_t rbytes = r->pack_desc_off + (size_t)r->g.kp_cap * 32;
  CUORB_CUDA_TRY(cudaMemcpyAsync(r->h_pack, r->d_pack, rbytes, cudaMemcpyDeviceToHost, r->stream));
  rc = cuorb_stereo_match_batch_async(l, r, mb, mbf);          // on l's stream, after r's ev_done
  if (rc != CUORB_OK) return rc;
  const size_t lbytes = l->pack_depth_off + (size_t)l->g.kp_cap * sizeof(float);
  CUORB_CUDA_TRY(cudaMemcpyAsync(l->h_pack, l->d_pack, lbytes, cudaMemcpyDeviceToHost, l->stream));
  CUORB_CUDA_TRY(cudaStreamSynchronize(r->stream));
  CUORB_CUDA_TRY(cudaStreamSynchronize(l->stream));
  cuorb_extractor* hs[2] = {l, r};
  cuorb_keypoint* ko[2] = {kps_left, kps_right};
  uint8_t* dd[2] = {desc_left, desc_right};
  int* no[2] = {n_left, n_right};
  for (int k = 0; k < 2; k++) {
    cuorb_extractor* e = hs[k];
    const int n1 = *(const int*)(e->h_pack + e->pack_n_off);
    *e->h_flags = *(const int*)e->h_pack;
    e->h_out_n[0] = n1; e->h_out_mono[0] = *(const int*)(e->h_pack + e->pack_mono_off);
    e->fetch_n = 1; e->fetch_capacity = capacity;
    *no[k] = n1;
    rc = check_flags(e);
    if (rc != CUORB_OK) return rc;
    if (n1 > capacity) { set_error("caller capacity smaller than the keypoint count"); return CUORB_E_CAPACITY; }
    if (n1 > 0) {
      memcpy(ko[k], e->h_pack + e->pack_kps_off, sizeof(cuorb_keypoint) * (size_t)n1);
      memcpy(dd[k], e->h_pack + e->pack_desc_off, 32 * (size_t)n1);
    }
  }
  const int nl = *n_left;
  if (nl > 0) {
    memcpy(u_right, l->h_pack + l->pack_ur_off, sizeof(float) * (size_t)nl);
    memcpy(depth, l->h_pack + l->pack_depth_off, sizeof(float) * (size_t)nl);
  }
  return l->h_out_mono[0];
}

int cuorb_stage_ms(cuorb_extractor* e, float* ms_out) {
  if (!e || !ms_out) { set_error("null argument"); return CUORB_E_BADARG; }
  if (!e->have_stage_times) { set_error("no profiled batch (call cuorb_set_profiling before extract)"); return CUORB_E_STATE; }
  CUORB_CUDA_TRY(cudaSetDevice(e->cfg.device));
  CUORB_CUDA_TRY(cudaStreamSynchronize(e->stream));
  for (int s = 0; s < CUORB_STAGE_COUNT; s++) {
    ms_out[s] = 0.f;
    if (s == CUORB_STAGE_STEREO && !e->stereo_timed) continue;
    CUORB_CUDA_TRY(cudaEventElapsedTime(&ms_out[s], e->ev_stage[s], e->ev_stage[s + 1]));
  }
  return CUORB_OK;
}

int cuorb_hamming256(const void* a, const void* b) {
  const uint64_t* pa = (const uint64_t*)a; const uint64_t* pb = (const uint64_t*)b;
  uint64_t x[4]; memcpy(x, pa, 32);
  uint64_t y[4]; memcpy(y, pb, 32);
  return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) +
         __builtin_popcountll(x[2] ^ y[2]) + __builtin_popcountll(x[3] ^ y[3]);
}

}  // extern "C"
