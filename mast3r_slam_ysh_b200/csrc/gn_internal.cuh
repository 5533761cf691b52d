// Internal layout + launcher prototypes shared by gn_hessian.cu / gn_solve.cu / gn_api.cu.
#pragma once
#include "m3r_common.cuh"

namespace m3r {

constexpr int kPoseDim = 7;          // sim(3) tangent (tau, omega, sigma)
constexpr int kAcc = 35;             // 28 lower-triangle entries of S = sum w J J^T  +  7 entries of g = sum w e J
constexpr int kAccStride = 36;       // padded stride of one tile's partial sums
constexpr int kTilePix = 8192;       // pixels of one edge compacted by one CTA of the packing pre-pass (32 per thread)
// Hessian kernel shape: two records per thread in packed arithmetic need ~70 accumulator registers, so the CTA is small
// and three of them share an SM (168 registers per thread); -DM3R_HESS_THREADS / -DM3R_HESS_CTAS time other shapes.
#ifndef M3R_HESS_THREADS
#define M3R_HESS_THREADS 128
#endif
#ifndef M3R_HESS_CTAS
#define M3R_HESS_CTAS 3
#endif
constexpr int kHessThreads = M3R_HESS_THREADS;
constexpr int kHessCtasPerSm = M3R_HESS_CTAS;
constexpr int kHessChunkAlign = 512;  // the Hessian kernel's per-CTA range of compacted records is a multiple of this
constexpr int kMaxHessCtas = 1024;   // upper bound on the persistent Hessian grid (2 CTAs per SM), sizes `partial`
constexpr int kBlk = 32;             // Cholesky block size
constexpr int kSmallMaxN = 224;      // largest system of the single-CTA shared-memory solver (gn_small.cuh)
constexpr int kBlockDoubles = 56;    // per-edge block: M (7x7 row-major) + v (7)

// device-resident solver state (ints)
enum StateSlot {
  kDone = 0, kIters = 1, kFailCount = 2, kFailNow = 3, kBadEdges = 4, kConverged = 5,
  kBarCount = 6, kBarGen = 7,  // grid barrier of the fused solver kernel
  kCFail = 8,                  // some point of some keyframe fails the confidence gate (set by pack_points)
  kPeerTimeout = 9,            // a peer rank never arrived at the in-kernel all-reduce (multi-GPU path)
  kSkyLanes = 10,              // > 0: the envelope (skyline) factorisation applies and needs this many lanes (gn_solve.cu)
  kNdSegs = 11,                // > 0: the envelope factorisation runs segmented over this many interiors (nd_plan.cuh)
  kNdSlot = 12,                // which of the planner's candidate plans it is
  kStateInts = 13
};

// Envelope ("skyline") plan of the assembled system, built once per solve by skyline_plan_kernel (gn_solve.cu) from the
// keyframe graph and consumed by the single-CTA envelope Cholesky in gn_fused.cu.  32-row blocks; fb[rb] = first column
// block of row block rb with a structural non-zero.  Row block rb is ACTIVE at elimination step kb when fb[rb] <= kb <=
// rb; an active row block owns a lane (< kSkyMaxLanes) from the step it joins until it has been eliminated itself.
constexpr int kSkyMaxLanes = 6;      // 21 resident 32x32 tiles of doubles = 183 KB of shared memory
constexpr int kSkyMaxBlk = 112;      // column blocks (n <= 3584: 512 keyframes, the reference's cap, frame.py:221)
constexpr int kSkySched = 8;         // ints per step: [m, new-member bitmask, act[0..5]] (act ascending, act[0] = kb)
struct SkyPlan {                     // int offsets inside the `sky` workspace region
  static __host__ __device__ int first(int) { return 0; }                          // [n_free] pose level (index prep)
  static __host__ __device__ int fb(int n_free) { return n_free; }                 // [kSkyMaxBlk]
  static __host__ __device__ int lane(int n_free) { return n_free + kSkyMaxBlk; }  // [kSkyMaxBlk]
  static __host__ __device__ int sched(int n_free) { return n_free + 2 * kSkyMaxBlk; }  // [kSkyMaxBlk][kSkySched]
  static __host__ __device__ int adj(int n_free) { return n_free + (2 + kSkySched) * kSkyMaxBlk; }  // [kSkyMaxBlk][4] bit cb of row rb:
                                                                     // tile (rb, cb), cb <= rb, has a structural non-zero
  static __host__ __device__ int total(int n_free) { return n_free + (6 + kSkySched) * kSkyMaxBlk + 8; }
};

// phases of the fused solver kernel (gn_fused.cu)
enum FusedPhase { M3R_PH_FINALIZE = 1, M3R_PH_ASSEMBLE = 2, M3R_PH_FACTOR = 4, M3R_PH_BACKSOLVE = 8, M3R_PH_ALL = 15 };

// Multi-GPU: the ranks' [H | b] buffers live in peer-mapped (symmetric) memory and the fused solver kernel sums them
// itself between assembly and factorisation (gn_fused.cu, peer_allreduce_phase) -- no NCCL call in the iteration.
constexpr int kMaxPeers = 8;
struct PeerInfo {
  int rank, world;
  unsigned long long epoch;              // strictly increasing per launch, the same on every rank
  double* sys[kMaxPeers];                // every rank's [H | b], addressable from this device
  unsigned long long* flags[kMaxPeers];  // every rank's arrival slots: flags[p][src] written by rank src
};

// Edge records produced once per solve by the packing pre-pass (iteration-invariant part of the per-pixel work).
// Only pixels that pass every pose-independent gate (valid_match & q > Q_thresh & c_i > C_thresh & c_j > C_thresh,
// gn_kernels.cu:953-957; calib: z_i > z_eps, :1368) get a record; the others contribute exactly zero in the reference
// (zero weight, :963-964) and are dropped.  Records of edge e are compacted, in pixel order, into
// rec[e*num_pts + 0 .. ecount[e]): 32 bytes holding everything the per-iteration kernel needs that does not depend on
// the poses (layout per residual kind: make_rec_a in gn_hessian.cu) -- one contiguous stream per edge, no gather.
struct __align__(32) GnRecord {
  float4 a, b;
};

struct GnLayout {
  int num_kf, num_pts, num_edges, n_free, n, ld, nblk, tiles;
  size_t off_rec, off_ie, off_je, off_io, off_jo, off_partial, off_blocks, off_ecount, off_eoff, off_tstate, off_inc_off, off_inc_list, off_inc_other, off_sys,
      off_lfac, off_pflag, off_sky, off_nd, off_dx, off_state, off_uniq_flag, off_all, off_K, total;
};

GnLayout gn_layout(int num_kf, int num_pts, int num_edges, int num_all_edges);

struct GnWs {
  GnLayout L;
  char* base;
  GnRecord* rec() const { return (GnRecord*)(base + L.off_rec); }
  int* ecount() const { return (int*)(base + L.off_ecount); }      // [E] records per edge
  long long* eoff() const { return (long long*)(base + L.off_eoff); }  // [E+1] exclusive prefix of ecount
  int* tstate() const { return (int*)(base + L.off_tstate); }      // [0] ticket, [1 + e*tiles + t] chained-scan state
  int* ie() const { return (int*)(base + L.off_ie); }
  int* je() const { return (int*)(base + L.off_je); }
  float* partial() const { return (float*)(base + L.off_partial); }
  double* blocks() const { return (double*)(base + L.off_blocks); }
  int* inc_off() const { return (int*)(base + L.off_inc_off); }
  int* inc_list() const { return (int*)(base + L.off_inc_list); }
  int* inc_other() const { return (int*)(base + L.off_inc_other); }  // per list entry: other end's free index, see gn_solve.cu
  double* sys() const { return (double*)(base + L.off_sys); }
  double* lfac() const { return (double*)(base + L.off_lfac); }
  int* pflag() const { return (int*)(base + L.off_pflag); }
  int* sky() const { return (int*)(base + L.off_sky); }  // SkyPlan, then NdPlan (nd_plan.cuh) when off_nd != 0
  double* ndbuf() const { return (double*)(base + L.off_nd); }  // scratch of the segmented envelope solver (gn_fused.cu)
  float* dx() const { return (float*)(base + L.off_dx); }
  int* state() const { return (int*)(base + L.off_state); }
  int* uniq_flag() const { return (int*)(base + L.off_uniq_flag); }
  long long* all_ids() const { return (long long*)(base + L.off_all); }
  float* Kdev() const { return (float*)(base + L.off_K); }
};

// gn_hessian.cu
int launch_pack_points(const GnWs& ws, const float* Xs, const float* Cs, float C_thresh, cudaStream_t st);
int launch_pack_edges(const GnWs& ws, const m3r_gn_params& p, const float* Xs, const float* Cs, const int64_t* idx, const uint8_t* valid,
                      const float* Q, cudaStream_t st);
int hessian_grid();
int launch_hessian(const GnWs& ws, const m3r_gn_params& p, const float* Twc, cudaStream_t st);
int launch_finalize_edges(const GnWs& ws, const float* Twc, cudaStream_t st);
int launch_export_blocks(const GnWs& ws, float* Hs, float* gs, cudaStream_t st);

// gn_solve.cu
int launch_index_prep(const GnWs& ws, const int64_t* local_ii, const int64_t* local_jj, const int64_t* all_ii,
                      const int64_t* all_jj, int num_all_edges, cudaStream_t st, cudaEvent_t* plan_join);
int join_solver_plans(cudaEvent_t plan_join, cudaStream_t st);
int launch_direct_edge_index(const GnWs& ws, const int64_t* ii_edge, const int64_t* jj_edge, cudaStream_t st);

// gn_fused.cu
int launch_fused(const GnWs& ws, float* Twc, float delta_thresh, int phases, cudaStream_t st,
                 const PeerInfo* peer = nullptr);
int launch_pose_retr(float* poses, const float* dx, int num_poses, int num_fix, cudaStream_t st);

}  // namespace m3r
