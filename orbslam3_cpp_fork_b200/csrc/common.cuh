// libcuorb internal definitions shared by all translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdlib>
#include <string>
#include <vector>

#include "../../include/cuorb.h"

namespace cuorb {

constexpr int EDGE_THRESHOLD  = 19;  // src/ORBextractor.cc:64
constexpr int HALF_PATCH_SIZE = 15;  // :63
constexpr int PATCH_SIZE      = 31;  // :62
constexpr int FAST_BORDER     = 16;  // minBorderX = EDGE_THRESHOLD - 3 (:761)
constexpr int ROW_LEAD        = 32;  // bytes before pixel x=0 in a padded row (19 border + 13 pad)
constexpr int MAX_LEVELS      = CUORB_MAX_LEVELS;

// Geometry of one pyramid level, identical for every image of a batch.  Computed on the host
// with the reference's float arithmetic (src/ORBextractor.cc:755-775, 1140-1146).
struct LevelGeom {
  int w, h;             // level size (sz at :1143)
  int pitch;            // bytes per row of the padded plane (multiple of 64)
  int bpitch;           // bytes per row of the blurred plane (multiple of 16)
  unsigned long long plane_off;  // byte offset of padded row 0 inside an image's pyramid block
  unsigned long long pix_off;    // byte offset of pixel (0,0)  (= plane_off + 19*pitch + ROW_LEAD)
  unsigned long long blur_off;   // byte offset inside an image's blurred block
  // FAST cell grid (:769-775)
  int nCols, nRows, wCell, hCell;
  int maxBorderX, maxBorderY;  // cols-16, rows-16
  int cell_base;               // index of this level's first cell in the flattened cell list
  // octree (:556-585)
  int   quota;   // mnFeaturesPerLevel[level]
  int   nIni;
  float hX;
  int   node_cap;              // node pool size for this level
  // candidate / selected-key storage (entries, per image)
  int cand_cap;
  unsigned long long cand_off; // offset (entries) inside an image's candidate block
  int key_cap;
  int key_off;                 // offset (entries) inside an image's selected-key block
  // resize tables (device offsets, in entries, into the table arrays)
  int xtab_off, ytab_off;
  float scale, inv_scale;      // mvScaleFactor / mvInvScaleFactor
  float kp_size;               // (float)(int)(31 * scale)   (:865)
};

struct ExtractorGeom {
  int       nlevels;
  int       ini_th, min_th;
  int       total_cells;
  int       max_cell_w, max_cell_h;
  int       kp_cap;                       // per-image output capacity (sum of key_cap)
  unsigned long long pyr_img_stride;      // bytes per image in the pyramid arena
  unsigned long long blur_img_stride;
  unsigned long long cand_img_stride;     // entries
  int       key_img_stride;               // entries
  LevelGeom lv[MAX_LEVELS];
};

// one entry of the horizontal / vertical resize table: source index + the two 11-bit weights
struct ResizeTap { short s, a0, a1, pad; };

// selected key of a level (after the quadtree): level coordinates incl. the 16-px border
struct LevelKey { unsigned short x, y; unsigned short score, pad; };

// packed FAST candidate: x:12 | y:12 | score:8, coordinates relative to (minBorderX,minBorderY)
__host__ __device__ inline unsigned pack_cand(int x, int y, int s) {
  return ((unsigned)x << 20) | ((unsigned)y << 8) | (unsigned)s;
}
__host__ __device__ inline int cand_x(unsigned c) { return (int)(c >> 20); }
__host__ __device__ inline int cand_y(unsigned c) { return (int)((c >> 8) & 0xFFF); }
__host__ __device__ inline int cand_s(unsigned c) { return (int)(c & 0xFF); }

// error plumbing
void        set_error(const std::string& msg);
extern std::atomic<unsigned long long> g_launch_count;
inline void count_launch(int n = 1) { g_launch_count.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

#define CUORB_CUDA_TRY(expr)                                                                  \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      ::cuorb::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));                 \
      return CUORB_E_CUDA;                                                                    \
    }                                                                                         \
  } while (0)

// ---- programmatic dependent launch (single-frame latency) ----
// A one-image extraction is a chain of ~13 kernels of 5-15 us each on one stream, so the gaps between them are a
// visible part of the call.  Every kernel of the chain starts with pdl_enter(): wait until the previous grid has
// completed and its writes are visible (a no-op when the kernel was not launched as a dependent), then let the NEXT
// kernel of the stream be scheduled already, so its CTAs sit at their own pdl_enter() when this grid drains.  The
// trigger comes after the wait, so at most one successor is resident at a time.  launch_k() adds the launch
// attribute when t_pdl is set (cuorb_api.cu sets it for batches of at most two images; for large batches resident,
// blocked successor CTAs would only take SM resources from the running kernel).
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
#endif
extern thread_local bool t_pdl;
// CUORB_PDL=0 turns every dependent-launch chain of the library into plain stream order (A/B and test hook)
inline bool pdl_allowed() { static const bool on = []() { const char* v = getenv("CUORB_PDL"); return !v || atoi(v) != 0; }(); return on; }
// A kernel whose predecessor in the stream is not a kernel of the chain (a copy, a memset, a cross-stream event wait)
// is launched with plain stream ordering: the attribute is only meant for kernel-after-kernel edges.
struct PdlPause { bool saved; PdlPause() : saved(t_pdl) { t_pdl = false; } ~PdlPause() { t_pdl = saved; } };
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = t_pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// device flags written by kernels (per handle)
enum { FLAG_CAND_OVERFLOW = 1, FLAG_NODE_OVERFLOW = 2, FLAG_KEY_OVERFLOW = 4 };

// ---- kernel launchers (one per stage; all asynchronous on `st`) ----
void build_resize_words(const ExtractorGeom& g, const std::vector<ResizeTap>& xtab, std::vector<uint4>& wt, int* off);
// launch plan of the staged resize kernel (k_pyr_resize_stage) for one level; tx == 0: not usable there
constexpr int PYR_MAX_COLBLOCKS = 4;
struct PyrPlan {
  int tx = 0, nb = 0;     // CTA width (threads = output words), CTAs per output row
  int max_rows = 0;       // source rows a strip can need
  int smem = 0;           // dynamic shared memory per CTA
  int xlo[PYR_MAX_COLBLOCKS] = {}, tw[PYR_MAX_COLBLOCKS] = {};   // per column block: first staged source byte (16-aligned), staged bytes per row
};
void build_pyr_plan(const ExtractorGeom& g, const std::vector<uint4>& wt, const int* wtab_off,
                    const std::vector<ResizeTap>& ytab, PyrPlan* plan);
void launch_pyramid(const ExtractorGeom& g, const ExtractorGeom* dg, const uint8_t* imgs, int w, int h,
                    size_t stride, size_t image_stride, int n_img, uint8_t* pyr, const ResizeTap* xtab,
                    const ResizeTap* ytab, const uint4* wtab, const int* wtab_off, const PyrPlan* plan, cudaStream_t st,
                    int* zero_counts = nullptr, int* zero_flags = nullptr);   // non-null: level 0 clears them (small batches)
void build_blur_tiles(const ExtractorGeom& g, std::vector<int4>& tiles);
void launch_blur(const ExtractorGeom* g, const ExtractorGeom* dg, int n_img, const int4* tiles,
                 int n_tiles, const uint8_t* pyr, uint8_t* blur, cudaStream_t st);
void build_fast_cells(const ExtractorGeom& g, std::vector<int4>& cells);
void launch_fast(const ExtractorGeom& g, const ExtractorGeom* dg, const int4* cells, int n_img, const uint8_t* pyr,
                 unsigned* cand, int* cand_count, int* flags, cudaStream_t st);
struct FastMaps;
int  build_fast_recs(const ExtractorGeom& g, std::vector<int4>& recs);
bool fast_persist_supported(const ExtractorGeom& g);
void launch_fast_persist(const ExtractorGeom& g, const FastMaps& maps, const int4* recs, int n_cells, int n_img,
                         unsigned* cand, int* cand_count, int* flags, int* work_counter, cudaStream_t st);
void launch_octree(const ExtractorGeom& g, const ExtractorGeom* dg, int n_img, const unsigned* cand,
                   unsigned* cand_scratch, const int* cand_count, LevelKey* keys, int* key_count,
                   int* flags, cudaStream_t st);
struct TmaMaps;
void launch_describe(const ExtractorGeom& g, const ExtractorGeom* dg, const TmaMaps& tmaps, int n_img, const uint8_t* pyr,
                     const uint8_t* blur, const LevelKey* keys, const int* key_count, int lap0,
                     int lap1, int* slot_map, cuorb_keypoint* out_kps, uint8_t* out_desc,
                     int* out_n, int* out_mono, cudaStream_t st);
int  init_describe_constants();

struct StereoArgs {
  const ExtractorGeom* dgL;
  const uint8_t *pyrL, *pyrR;
  const cuorb_keypoint *kpsL, *kpsR;
  const uint8_t *descL, *descR;
  const int *nL, *nR;
  int kp_cap;      // stride (entries) between images in kps/desc arrays
  int n_img;
  float mb, mbf;
  float *u_right, *depth;
  int *best_idx_r, *best_dist, *sad;
  uint4* rtab;   // right keypoints bucketed by row band: {minr | maxr<<13 | octave<<26, x bits, iR, 0}
  int*   rstart; // per image STEREO_MAX_BUCKETS + 1 bucket starts
};
constexpr int STEREO_MAX_BUCKETS = 512;  // 8-row buckets: level-0 height below 4096
void launch_stereo(const ExtractorGeom& g, const StereoArgs& a, cudaStream_t st);

// pre-extraction image ops (preproc.cu): src = n images of sh x sw x channels, dst = gray dh x dw
struct PreArgs {
  const uint8_t* src; size_t sstride, simg_stride; int sw, sh;
  uint8_t* dst; size_t dstride, dimg_stride; int dw, dh;
  const float* mapx; const float* mapy;                                   // mode 1 (remap), dh x dw each
  const int* xofs; const short* xa; const int* yofs; const short* yb;     // mode 2 (resize) tables
  int area2;                                                              // mode 2: exact 2x2 decimation
  int ir, ib;                                                             // channel index of R and B
};
void launch_preproc(const PreArgs& a, int channels, int mode, int n_img, cudaStream_t st);

}  // namespace cuorb
