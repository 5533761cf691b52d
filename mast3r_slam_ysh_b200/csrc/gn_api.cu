// C ABI of the Gauss-Newton path (include/m3r_backends.h): workspace layout, the fused on-device GN loop, the
// step-wise variant for edge-sharded multi-GPU runs, diagnostics and host-buffer wrappers.
// Host control flow replaces gauss_newton_{points,rays,calib}_cuda (gn_kernels.cu:725-811, :1140-1228, :1546-1637).
#include <stdarg.h>
#include <string.h>

#include <atomic>

#include "gn_internal.cuh"
#include "nd_plan.cuh"

namespace m3r {

// ---- error / counters (shared by all translation units) ---------------------------------------------------------
static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

int current_device_slot() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) dev = 0;
  return dev < kMaxDevices ? dev : kMaxDevices - 1;
}

int sm_count() {
  static std::atomic<int> cached[kMaxDevices];
  const int d = current_device_slot();
  int n = cached[d].load(std::memory_order_relaxed);
  if (n == 0) {
    int dev = 0;
    n = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
    cached[d].store(n, std::memory_order_relaxed);
  }
  return n;
}

int keep_async_pool() {
  static thread_local int ready_dev = -1;
  int dev = 0;
  M3R_CUDA(cudaGetDevice(&dev));
  if (ready_dev != dev) {
    cudaMemPool_t pool;
    M3R_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
    uint64_t thr = UINT64_MAX;
    M3R_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
    ready_dev = dev;
  }
  return M3R_OK;
}

// ---- workspace layout ---------------------------------------------------------------------------------------------
GnLayout gn_layout(int num_kf, int num_pts, int num_edges, int num_all_edges) {
  GnLayout L;
  memset(&L, 0, sizeof(L));
  L.num_kf = num_kf;
  L.num_pts = num_pts;
  L.num_edges = num_edges;
  L.n_free = num_kf > 0 ? num_kf - 1 : 0;
  L.n = kPoseDim * L.n_free;
  L.ld = (int)align_up((size_t)(L.n > 0 ? L.n : 1), 8);
  L.nblk = ceil_div(L.n, kBlk);
  L.tiles = num_pts > 0 ? ceil_div(num_pts, kTilePix) : 0;
  size_t o = 0;
  auto take = [&](size_t bytes) {
    size_t at = o;
    o = align_up(o + bytes, 256);
    return at;
  };
  const size_t E = (size_t)num_edges, Ea = (size_t)(num_all_edges > num_edges ? num_all_edges : num_edges);
  L.off_state = take(sizeof(int) * kStateInts);
  L.off_K = take(sizeof(float) * 16);
  L.off_dx = take(sizeof(float) * (size_t)(L.n + 8));
  L.off_ie = take(sizeof(int) * (E + 1));
  L.off_je = take(sizeof(int) * (E + 1));
  L.off_io = take(8);
  L.off_jo = take(8);
  L.off_inc_off = take(sizeof(int) * (size_t)(L.n_free + 2));
  L.off_inc_list = take(sizeof(int) * (2 * E + 2));
  L.off_inc_other = take(sizeof(int) * (2 * E + 2));
  L.off_all = take(sizeof(long long) * (2 * Ea + 2));
  L.off_uniq_flag = take(sizeof(int) * (2 * Ea + 2));
  L.off_blocks = take(sizeof(double) * kBlockDoubles * (E + 1));
  L.off_partial = take(sizeof(float) * kAccStride * (E + kMaxHessCtas + 1));
  L.off_ecount = take(sizeof(int) * (E + 1));
  L.off_eoff = take(sizeof(long long) * (E + 2));
  L.off_tstate = take(sizeof(int) * (E * (size_t)L.tiles + 2));
  L.off_sys = take(sizeof(double) * (size_t)L.ld * (size_t)(L.n + 1));
  L.off_lfac = take(sizeof(double) * kBlk * kBlk * (size_t)(L.nblk + 1));
  L.off_pflag = take(sizeof(int) * (size_t)(L.n / 64 + 4) * (size_t)(L.n / 64 + 5));  // pflag + pmask rows
  // the envelope solver's plan; systems it can take (gn_fused.cu) also get the segmented schedule and its scratch
  const bool nd = L.n > kSmallMaxN && L.nblk <= kSkyMaxBlk;
  L.off_sky = take(sizeof(int) * (size_t)(SkyPlan::total(L.n_free) + (nd ? kNdCandP * kNdCandSw * NdPlan::total + 8 : 0)));
  L.off_nd = nd ? take(sizeof(double) * NdBuf::total) : 0;
  L.off_rec = take(sizeof(GnRecord) * E * (size_t)num_pts);
  L.total = o;
  return L;
}

static int check_sizes(int num_kf, int num_pts, int num_edges) {
  M3R_REQUIRE(num_kf >= 0 && num_pts >= 0 && num_edges >= 0, "gauss_newton: negative size");
  M3R_REQUIRE((long long)num_kf * num_pts < (1LL << 40), "gauss_newton: problem too large");
  M3R_REQUIRE((long long)num_edges * ceil_div(num_pts > 0 ? num_pts : 1, kTilePix) < (1LL << 31),
              "gauss_newton: too many edge tiles");
  return M3R_OK;
}

static int check_params(const m3r_gn_params* p) {
  M3R_REQUIRE(p != nullptr, "gauss_newton: params is NULL");
  M3R_REQUIRE(p->kind >= M3R_GN_POINTS && p->kind <= M3R_GN_CALIB, "gauss_newton: unknown kind %d", p->kind);
  M3R_REQUIRE(p->robust == M3R_ROBUST_HUBER || p->robust == M3R_ROBUST_TUKEY, "gauss_newton: unknown robust kernel %d",
              p->robust);
  if (p->kind == M3R_GN_CALIB) {
    M3R_REQUIRE(p->K != nullptr, "gauss_newton_calib: K is NULL");
    M3R_REQUIRE(p->width > 0 && p->height > 0, "gauss_newton_calib: bad image size %dx%d", p->width, p->height);
  }
  return M3R_OK;
}

static int bind_ws(GnWs* ws, int num_kf, int num_pts, int num_edges, int num_all_edges, void* workspace,
                   size_t workspace_bytes) {
  ws->L = gn_layout(num_kf, num_pts, num_edges, num_all_edges);
  M3R_REQUIRE(workspace != nullptr, "gauss_newton: workspace is NULL");
  M3R_REQUIRE(((uintptr_t)workspace & 255) == 0, "gauss_newton: workspace must be 256-byte aligned");
  if (workspace_bytes < ws->L.total) {
    set_error("gauss_newton: workspace too small (%zu < %zu bytes)", workspace_bytes, ws->L.total);
    return M3R_ERR_WORKSPACE_TOO_SMALL;
  }
  ws->base = (char*)workspace;
  return M3R_OK;
}

__global__ void finish_kernel(const float* __restrict__ dx, const int* __restrict__ state, float* __restrict__ dx_out,
                              int32_t* __restrict__ info_out, int n) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (dx_out != nullptr && t < n) dx_out[t] = dx[t];
  if (info_out != nullptr && t < 4) {
    const int v = t == 0   ? state[kIters]
                  : t == 1 ? state[kConverged]
                  : t == 2 ? state[kFailCount]
                           : ((state[kBadEdges] & 0xfffff) | ((state[kNdSegs] & 15) << 20) | ((state[kSkyLanes] & 15) << 24) |
                              (state[kPeerTimeout] ? 0x40000000 : 0));
    info_out[t] = v;
  }
}

__global__ void zero_f32_kernel(float* p, int n) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) p[t] = 0.0f;
}

static int gn_begin_impl(GnWs& ws, const m3r_gn_params* p, const float* Xs, const float* Cs, const int64_t* local_ii,
                         const int64_t* local_jj, const int64_t* idx, const uint8_t* valid, const float* Q,
                         const int64_t* all_ii, const int64_t* all_jj, int num_all_edges, cudaStream_t st) {
  int rc;
  cudaEvent_t plan_join = nullptr;  // the solver plans run next to the packing pre-pass (gn_solve.cu)
  if ((rc = launch_index_prep(ws, local_ii, local_jj, all_ii, all_jj, num_all_edges, st, &plan_join)) != M3R_OK) return rc;
  if (p->kind == M3R_GN_CALIB)
    M3R_CUDA(cudaMemcpyAsync(ws.Kdev(), p->K, sizeof(float) * 9, cudaMemcpyDeviceToDevice, st));
  // the full system buffer is cleared once so the (never read) upper triangle stays finite for the all-reduce
  M3R_CUDA(cudaMemsetAsync(ws.sys(), 0, sizeof(double) * (size_t)ws.L.ld * (size_t)(ws.L.n + 1), st));
  if (ws.L.n > 0) {
    zero_f32_kernel<<<ceil_div(ws.L.n, 256), 256, 0, st>>>(ws.dx(), ws.L.n);
    M3R_LAUNCH_CHECK();
  }
  if ((rc = launch_pack_points(ws, Xs, Cs, p->C_thresh, st)) == M3R_OK)
    rc = launch_pack_edges(ws, *p, Xs, Cs, idx, valid, Q, st);
  const int rcj = join_solver_plans(plan_join, st);  // also on the error path: the event must not leak
  return rc != M3R_OK ? rc : rcj;
}

// one GN iteration = the Hessian kernel + ONE fused solver launch; the sharded path splits the latter at the all-reduce
static int gn_accumulate_impl(GnWs& ws, const m3r_gn_params* p, const float* Twc, cudaStream_t st) {
  int rc;
  if ((rc = launch_hessian(ws, *p, Twc, st)) != M3R_OK) return rc;
  return launch_fused(ws, const_cast<float*>(Twc), p->delta_thresh, M3R_PH_FINALIZE | M3R_PH_ASSEMBLE, st);
}

static int gn_solve_impl(GnWs& ws, const m3r_gn_params* p, float* Twc, cudaStream_t st) {
  return launch_fused(ws, Twc, p->delta_thresh, M3R_PH_FACTOR | M3R_PH_BACKSOLVE, st);
}

static int gn_iterate_impl(GnWs& ws, const m3r_gn_params* p, float* Twc, cudaStream_t st) {
  int rc;
  if ((rc = launch_hessian(ws, *p, Twc, st)) != M3R_OK) return rc;
  return launch_fused(ws, Twc, p->delta_thresh, M3R_PH_ALL, st);
}

// multi-GPU iteration with the in-kernel peer all-reduce: Hessian kernel + ONE fused launch, as on a single GPU
static int gn_iterate_peer_impl(GnWs& ws, const m3r_gn_params* p, float* Twc, const PeerInfo& peer, cudaStream_t st) {
  int rc;
  if ((rc = launch_hessian(ws, *p, Twc, st)) != M3R_OK) return rc;
  return launch_fused(ws, Twc, p->delta_thresh, M3R_PH_ALL, st, &peer);
}

static int gn_end_impl(GnWs& ws, float* dx_out, int32_t* info_out, cudaStream_t st) {
  const int n = ws.L.n;
  const int cnt = n > 4 ? n : 4;
  finish_kernel<<<ceil_div(cnt, 256), 256, 0, st>>>(ws.dx(), ws.state(), dx_out, info_out, n);
  M3R_LAUNCH_CHECK();
  return M3R_OK;
}

}  // namespace m3r

using namespace m3r;

extern "C" int m3r_abi_version(void) { return M3R_ABI_VERSION; }
extern "C" const char* m3r_last_error_string(void) { return m3r::g_err; }
extern "C" uint64_t m3r_launch_count(void) { return m3r::g_launches.load(std::memory_order_relaxed); }

// the "all ranks' edges" scratch is sized for up to 16x the local edge count (uneven 8-way splits) or 4096 ids
static int max_all_edges(int num_edges) {
  const long long ea = (long long)num_edges * 16 + 64 > 4096 ? (long long)num_edges * 16 + 64 : 4096;
  return (int)ea;
}

extern "C" size_t m3r_gn_workspace_bytes(int num_kf, int num_pts, int num_edges) {
  if (num_kf < 0 || num_pts < 0 || num_edges < 0) return 0;
  return gn_layout(num_kf, num_pts, num_edges, max_all_edges(num_edges)).total;
}

extern "C" int m3r_gn_begin(const m3r_gn_params* params, const float* Twc, const float* Xs, const float* Cs,
                            const int64_t* local_ii, const int64_t* local_jj, const int64_t* idx_ii2jj,
                            const uint8_t* valid_match, const float* Q, int num_kf, int num_pts, int num_local_edges,
                            const int64_t* all_ii, const int64_t* all_jj, int num_all_edges, void* workspace,
                            size_t workspace_bytes, void* stream) {
  (void)Twc;
  int rc;
  if ((rc = check_params(params)) != M3R_OK) return rc;
  if ((rc = check_sizes(num_kf, num_pts, num_local_edges)) != M3R_OK) return rc;
  M3R_REQUIRE(num_all_edges >= num_local_edges && num_all_edges <= max_all_edges(num_local_edges),
              "gn_begin: num_all_edges=%d out of range for %d local edges", num_all_edges, num_local_edges);
  if (num_local_edges > 0 && num_pts > 0)
    M3R_REQUIRE(Xs && Cs && local_ii && local_jj && idx_ii2jj && valid_match && Q, "gn_begin: null pointer");
  if (num_all_edges > 0) M3R_REQUIRE(all_ii && all_jj, "gn_begin: null all_ii/all_jj");
  GnWs ws;
  if ((rc = bind_ws(&ws, num_kf, num_pts, num_local_edges, max_all_edges(num_local_edges), workspace,
                    workspace_bytes)) != M3R_OK)
    return rc;
  return gn_begin_impl(ws, params, Xs, Cs, local_ii, local_jj, idx_ii2jj, valid_match, Q, all_ii, all_jj,
                       num_all_edges, (cudaStream_t)stream);
}

extern "C" int m3r_gn_accumulate(const m3r_gn_params* params, const float* Twc, int num_kf, int num_pts,
                                 int num_local_edges, void* workspace, void* stream) {
  int rc;
  if ((rc = check_params(params)) != M3R_OK) return rc;
  GnWs ws;
  if ((rc = bind_ws(&ws, num_kf, num_pts, num_local_edges, max_all_edges(num_local_edges), workspace, (size_t)-1)) !=
      M3R_OK)
    return rc;
  return gn_accumulate_impl(ws, params, Twc, (cudaStream_t)stream);
}

extern "C" int m3r_gn_hessian_pass(const m3r_gn_params* params, const float* Twc, int num_kf, int num_pts,
                                   int num_local_edges, void* workspace, void* stream) {
  int rc;
  if ((rc = check_params(params)) != M3R_OK) return rc;
  GnWs ws;
  if ((rc = bind_ws(&ws, num_kf, num_pts, num_local_edges, max_all_edges(num_local_edges), workspace, (size_t)-1)) !=
      M3R_OK)
    return rc;
  return launch_hessian(ws, *params, Twc, (cudaStream_t)stream);
}

extern "C" int m3r_gn_iterate_peer(const m3r_gn_params* params, float* Twc, int num_kf, int num_pts,
                                   int num_local_edges, void* workspace, int rank, int world,
                                   const uint64_t* peer_sys_ptrs, const uint64_t* peer_flag_ptrs, uint64_t epoch,
                                   void* stream) {
  int rc;
  if ((rc = check_params(params)) != M3R_OK) return rc;
  M3R_REQUIRE(world >= 2 && world <= kMaxPeers && rank >= 0 && rank < world, "gn_iterate_peer: bad rank/world %d/%d",
              rank, world);
  M3R_REQUIRE(peer_sys_ptrs && peer_flag_ptrs && epoch > 0, "gn_iterate_peer: null peer table or zero epoch");
  GnWs ws;
  if ((rc = bind_ws(&ws, num_kf, num_pts, num_local_edges, max_all_edges(num_local_edges), workspace, (size_t)-1)) !=
      M3R_OK)
    return rc;
  PeerInfo peer = PeerInfo();
  peer.rank = rank;
  peer.world = world;
  peer.epoch = epoch;
  for (int p = 0; p < world; p++) {
    M3R_REQUIRE(peer_sys_ptrs[p] != 0 && (peer_sys_ptrs[p] & 15) == 0 && peer_flag_ptrs[p] != 0 &&
                    (peer_flag_ptrs[p] & 7) == 0,
                "gn_iterate_peer: peer %d pointers must be non-null and aligned", p);
    peer.sys[p] = reinterpret_cast<double*>(peer_sys_ptrs[p]);
    peer.flags[p] = reinterpret_cast<unsigned long long*>(peer_flag_ptrs[p]);
  }
  return gn_iterate_peer_impl(ws, params, Twc, peer, (cudaStream_t)stream);
}

extern "C" double* m3r_gn_system_ptr(void* workspace, int num_kf, int num_pts, int num_local_edges) {
  if (!workspace) return nullptr;
  GnLayout L = gn_layout(num_kf, num_pts, num_local_edges, max_all_edges(num_local_edges));
  return (double*)((char*)workspace + L.off_sys);
}

extern "C" size_t m3r_gn_system_count(int num_kf) {
  GnLayout L = gn_layout(num_kf, 0, 0, 0);
  return (size_t)L.ld * (size_t)(L.n + 1);
}

extern "C" long long m3r_gn_record_count(const void* workspace, int num_kf, int num_pts, int num_local_edges) {
  if (!workspace) return -1;
  GnLayout L = gn_layout(num_kf, num_pts, num_local_edges, max_all_edges(num_local_edges));
  long long v = 0;
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  if (cudaMemcpy(&v, (const char*)workspace + L.off_eoff + sizeof(long long) * (size_t)num_local_edges, sizeof(long long),
                 cudaMemcpyDeviceToHost) != cudaSuccess)
    return -1;
  return v;
}

extern "C" int m3r_gn_solve_retract(const m3r_gn_params* params, float* Twc, int num_kf, int num_pts,
                                    int num_local_edges, void* workspace, void* stream) {
  int rc;
  if ((rc = check_params(params)) != M3R_OK) return rc;
  GnWs ws;
  if ((rc = bind_ws(&ws, num_kf, num_pts, num_local_edges, max_all_edges(num_local_edges), workspace, (size_t)-1)) !=
      M3R_OK)
    return rc;
  return gn_solve_impl(ws, params, Twc, (cudaStream_t)stream);
}

extern "C" int m3r_gn_end(int num_kf, int num_pts, int num_local_edges, float* dx_out, int32_t* info_out,
                          void* workspace, void* stream) {
  int rc;
  GnWs ws;
  if ((rc = bind_ws(&ws, num_kf, num_pts, num_local_edges, max_all_edges(num_local_edges), workspace, (size_t)-1)) !=
      M3R_OK)
    return rc;
  return gn_end_impl(ws, dx_out, info_out, (cudaStream_t)stream);
}

extern "C" int m3r_gauss_newton(const m3r_gn_params* params, float* Twc, const float* Xs, const float* Cs,
                                const int64_t* ii, const int64_t* jj, const int64_t* idx_ii2jj,
                                const uint8_t* valid_match, const float* Q, int num_kf, int num_pts, int num_edges,
                                float* dx_out, int32_t* info_out, void* workspace, size_t workspace_bytes,
                                void* stream) {
  int rc;
  if ((rc = check_params(params)) != M3R_OK) return rc;
  if ((rc = check_sizes(num_kf, num_pts, num_edges)) != M3R_OK) return rc;
  M3R_REQUIRE(params->max_iter >= 0, "gauss_newton: max_iter < 0");
  if (num_kf > 0) M3R_REQUIRE(Twc != nullptr, "gauss_newton: Twc is NULL");
  if (num_edges > 0 && num_pts > 0)
    M3R_REQUIRE(Xs && Cs && ii && jj && idx_ii2jj && valid_match && Q, "gauss_newton: null pointer");
  GnWs ws;
  if ((rc = bind_ws(&ws, num_kf, num_pts, num_edges, max_all_edges(num_edges), workspace, workspace_bytes)) != M3R_OK)
    return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if ((rc = gn_begin_impl(ws, params, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, ii, jj, num_edges, st)) != M3R_OK)
    return rc;
  for (int it = 0; it < params->max_iter; it++) {
    if ((rc = gn_iterate_impl(ws, params, Twc, st)) != M3R_OK) return rc;
  }
  return gn_end_impl(ws, dx_out, info_out, st);
}

extern "C" int m3r_gn_edge_blocks(const m3r_gn_params* params, const float* Twc, const float* Xs, const float* Cs,
                                  const int64_t* ii_edge, const int64_t* jj_edge, const int64_t* idx_ii2jj,
                                  const uint8_t* valid_match, const float* Q, int num_kf, int num_pts, int num_edges,
                                  float* Hs, float* gs, void* workspace, size_t workspace_bytes, void* stream) {
  int rc;
  if ((rc = check_params(params)) != M3R_OK) return rc;
  if ((rc = check_sizes(num_kf, num_pts, num_edges)) != M3R_OK) return rc;
  M3R_REQUIRE(Hs && gs, "gn_edge_blocks: null output");
  GnWs ws;
  if ((rc = bind_ws(&ws, num_kf, num_pts, num_edges, max_all_edges(num_edges), workspace, workspace_bytes)) != M3R_OK)
    return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // ii_edge/jj_edge already index Twc/Xs (the reference's `ii_edge`), no unique/searchsorted here
  if ((rc = launch_direct_edge_index(ws, ii_edge, jj_edge, st)) != M3R_OK) return rc;
  if (params->kind == M3R_GN_CALIB)
    M3R_CUDA(cudaMemcpyAsync(ws.Kdev(), params->K, sizeof(float) * 9, cudaMemcpyDeviceToDevice, st));
  if ((rc = launch_pack_points(ws, Xs, Cs, params->C_thresh, st)) != M3R_OK) return rc;
  if ((rc = launch_pack_edges(ws, *params, Xs, Cs, idx_ii2jj, valid_match, Q, st)) != M3R_OK) return rc;
  if ((rc = launch_hessian(ws, *params, Twc, st)) != M3R_OK) return rc;
  if ((rc = launch_finalize_edges(ws, Twc, st)) != M3R_OK) return rc;
  return launch_export_blocks(ws, Hs, gs, st);
}

extern "C" int m3r_pose_retr(float* poses, const float* dx, int num_poses, int num_fix, void* stream) {
  M3R_REQUIRE(num_poses >= 0 && num_fix >= 0, "pose_retr: negative size");
  if (num_poses - num_fix <= 0) return M3R_OK;
  M3R_REQUIRE(poses && dx, "pose_retr: null pointer");
  return launch_pose_retr(poses, dx, num_poses, num_fix, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------------- host wrappers
namespace {
struct DevBuf {
  void* p = nullptr;
  ~DevBuf() {
    if (p) cudaFree(p);
  }
  int alloc(size_t bytes) { return (int)cudaMalloc(&p, bytes > 0 ? bytes : 1); }
};
}  // namespace

#define M3R_H2D(dst, src, bytes) M3R_CUDA(cudaMemcpyAsync((dst).p, (src), (bytes), cudaMemcpyHostToDevice, 0))

extern "C" int m3r_iter_proj_host(const float* rays, const float* pts, const float* p_init, float* p_new,
                                  uint8_t* converged, int b, int h, int w, int n, int max_iter, float lambda_init,
                                  float cost_thresh) {
  M3R_REQUIRE(b >= 0 && n >= 0 && h >= 0 && w >= 0, "iter_proj_host: negative size");
  const size_t img = sizeof(float) * 9 * (size_t)b * h * w, np = (size_t)b * n;
  if (np == 0) return M3R_OK;
  DevBuf d_r, d_p, d_i, d_o, d_c;
  M3R_CUDA((cudaError_t)d_r.alloc(img));
  M3R_CUDA((cudaError_t)d_p.alloc(np * 12));
  M3R_CUDA((cudaError_t)d_i.alloc(np * 8));
  M3R_CUDA((cudaError_t)d_o.alloc(np * 8));
  M3R_CUDA((cudaError_t)d_c.alloc(np));
  M3R_H2D(d_r, rays, img);
  M3R_H2D(d_p, pts, np * 12);
  M3R_H2D(d_i, p_init, np * 8);
  int rc = m3r_iter_proj((const float*)d_r.p, (const float*)d_p.p, (const float*)d_i.p, (float*)d_o.p, (uint8_t*)d_c.p,
                         b, h, w, n, max_iter, lambda_init, cost_thresh, nullptr);
  if (rc != M3R_OK) return rc;
  M3R_CUDA(cudaMemcpyAsync(p_new, d_o.p, np * 8, cudaMemcpyDeviceToHost, 0));
  M3R_CUDA(cudaMemcpyAsync(converged, d_c.p, np, cudaMemcpyDeviceToHost, 0));
  M3R_CUDA(cudaStreamSynchronize(0));
  return M3R_OK;
}

extern "C" int m3r_refine_matches_f16_host(const uint16_t* D11, const uint16_t* D21, const int64_t* p1, int64_t* p1_new,
                                           int b, int h, int w, int n, int fdim, int radius, int dilation_max) {
  M3R_REQUIRE(b >= 0 && n >= 0 && h >= 0 && w >= 0 && fdim >= 0, "refine_matches_host: negative size");
  const size_t np = (size_t)b * n;
  if (np == 0) return M3R_OK;
  DevBuf d_a, d_b, d_p, d_o;
  M3R_CUDA((cudaError_t)d_a.alloc(2 * (size_t)b * h * w * fdim));
  M3R_CUDA((cudaError_t)d_b.alloc(2 * np * fdim));
  M3R_CUDA((cudaError_t)d_p.alloc(np * 16));
  M3R_CUDA((cudaError_t)d_o.alloc(np * 16));
  M3R_H2D(d_a, D11, 2 * (size_t)b * h * w * fdim);
  M3R_H2D(d_b, D21, 2 * np * fdim);
  M3R_H2D(d_p, p1, np * 16);
  int rc = m3r_refine_matches_f16((const uint16_t*)d_a.p, (const uint16_t*)d_b.p, (const int64_t*)d_p.p,
                                  (int64_t*)d_o.p, b, h, w, n, fdim, radius, dilation_max, nullptr);
  if (rc != M3R_OK) return rc;
  M3R_CUDA(cudaMemcpyAsync(p1_new, d_o.p, np * 16, cudaMemcpyDeviceToHost, 0));
  M3R_CUDA(cudaStreamSynchronize(0));
  return M3R_OK;
}

extern "C" int m3r_gauss_newton_host(const m3r_gn_params* params, float* Twc, const float* Xs, const float* Cs,
                                     const int64_t* ii, const int64_t* jj, const int64_t* idx, const uint8_t* valid,
                                     const float* Q, int num_kf, int num_pts, int num_edges, float* dx_out,
                                     int32_t* info_out) {
  int rc;
  if ((rc = check_params(params)) != M3R_OK) return rc;
  if ((rc = check_sizes(num_kf, num_pts, num_edges)) != M3R_OK) return rc;
  const size_t KN = (size_t)num_kf * num_pts, EN = (size_t)num_edges * num_pts;
  DevBuf d_T, d_X, d_C, d_ii, d_jj, d_idx, d_v, d_Q, d_K, d_dx, d_info, d_ws;
  const size_t wsb = m3r_gn_workspace_bytes(num_kf, num_pts, num_edges);
  M3R_CUDA((cudaError_t)d_T.alloc(sizeof(float) * 8 * (size_t)num_kf));
  M3R_CUDA((cudaError_t)d_X.alloc(sizeof(float) * 3 * KN));
  M3R_CUDA((cudaError_t)d_C.alloc(sizeof(float) * KN));
  M3R_CUDA((cudaError_t)d_ii.alloc(8 * (size_t)num_edges));
  M3R_CUDA((cudaError_t)d_jj.alloc(8 * (size_t)num_edges));
  M3R_CUDA((cudaError_t)d_idx.alloc(8 * EN));
  M3R_CUDA((cudaError_t)d_v.alloc(EN));
  M3R_CUDA((cudaError_t)d_Q.alloc(4 * EN));
  M3R_CUDA((cudaError_t)d_K.alloc(64));
  M3R_CUDA((cudaError_t)d_dx.alloc(sizeof(float) * 7 * (size_t)(num_kf > 0 ? num_kf : 1)));
  M3R_CUDA((cudaError_t)d_info.alloc(16));
  M3R_CUDA((cudaError_t)d_ws.alloc(wsb));
  M3R_H2D(d_T, Twc, sizeof(float) * 8 * (size_t)num_kf);
  M3R_H2D(d_X, Xs, sizeof(float) * 3 * KN);
  M3R_H2D(d_C, Cs, sizeof(float) * KN);
  M3R_H2D(d_ii, ii, 8 * (size_t)num_edges);
  M3R_H2D(d_jj, jj, 8 * (size_t)num_edges);
  M3R_H2D(d_idx, idx, 8 * EN);
  M3R_H2D(d_v, valid, EN);
  M3R_H2D(d_Q, Q, 4 * EN);
  m3r_gn_params p = *params;
  if (p.kind == M3R_GN_CALIB) {
    M3R_H2D(d_K, params->K, sizeof(float) * 9);
    p.K = (const float*)d_K.p;
  }
  rc = m3r_gauss_newton(&p, (float*)d_T.p, (const float*)d_X.p, (const float*)d_C.p, (const int64_t*)d_ii.p,
                        (const int64_t*)d_jj.p, (const int64_t*)d_idx.p, (const uint8_t*)d_v.p, (const float*)d_Q.p,
                        num_kf, num_pts, num_edges, (float*)d_dx.p, (int32_t*)d_info.p, d_ws.p, wsb, nullptr);
  if (rc != M3R_OK) return rc;
  M3R_CUDA(cudaMemcpyAsync(Twc, d_T.p, sizeof(float) * 8 * (size_t)num_kf, cudaMemcpyDeviceToHost, 0));
  if (dx_out && num_kf > 1)
    M3R_CUDA(cudaMemcpyAsync(dx_out, d_dx.p, sizeof(float) * 7 * (size_t)(num_kf - 1), cudaMemcpyDeviceToHost, 0));
  if (info_out) M3R_CUDA(cudaMemcpyAsync(info_out, d_info.p, 16, cudaMemcpyDeviceToHost, 0));
  M3R_CUDA(cudaStreamSynchronize(0));
  return M3R_OK;
}
