// Frame tracker pose solve on the device for sm_100a: the whole Gauss-Newton loop of
//   FrameTracker.opt_pose_ray_dist_sim3 (mast3r_slam/tracker.py:173-214) and
//   FrameTracker.opt_pose_calib_sim3    (mast3r_slam/tracker.py:216-266), with FrameTracker.solve (:156-171),
//   geometry.act_Sim3 / point_to_ray_dist / project_calib (geometry.py:17-34, :45-52, :63-108),
//   nonlinear_optimizer.huber / check_convergence (nonlinear_optimizer.py:5-33)
// in ONE persistent cooperative kernel.  The reference runs <= 50 iterations of ~40 eager torch launches each with a
// .item() host sync per iteration; here a launch does
//   load   : each CTA takes 128-point tiles (strided over the grid, so gated-out image regions are spread evenly),
//            drops the points whose weight is zero (valid_match / C / Q gates, tracker.py:59-64) and keeps the
//            survivors in SHARED MEMORY for the whole solve (32 bytes per point: the frame point + sqrt(Q), and the
//            pre-computed keyframe ray/distance resp. the pixel/log-depth measurement);
//   iterate: per point the 4 (rays) or 3 (calib) weighted residual rows with their analytically sparse 7-column
//            Jacobians accumulate H (28) + g (7) + cost in fp32 registers; lanes reduce with a transposing butterfly,
//            warps through shared memory in fp64; the CTA publishes 36 doubles + a release flag; EVERY CTA then sums all
//            CTAs' partials in the same fixed order, solves the 7x7 system in fp64, retracts T_CkCf and evaluates the
//            convergence test redundantly -- bit-identical on all CTAs, so one all-to-all exchange per iteration is the only
//            grid-wide synchronisation and no broadcast of the pose is needed;
//   finish : T_WCf = T_WCk * T_CkCf (tracker.py:212), iteration count, final cost, Cholesky status.
// Points with zero weight contribute exactly 0 here; in the reference 0 * (inf or NaN residual) poisons H and the
// frame is skipped ("Cholesky failed"), which only happens for degenerate points (X = 0, or z = 0 in the calib case).
#include <cooperative_groups.h>

#include "gn_sim3.cuh"

namespace m3r {

constexpr int kTrkThreads = 512;
constexpr int kTrkWarps = kTrkThreads / 32;
constexpr int kTrkTile = 128;  // points per load tile
constexpr int kTrkTilesPerRound = kTrkThreads / kTrkTile;
constexpr int kTrkVals = 36;      // 28 (lower triangle of H) + 7 (J^T W r) + 1 (cost)
constexpr int kTrkMaxCtas = 160;  // >= SMs of a B200 (148); bounds the unrolled cross-CTA sum
constexpr int kTrkPerWarp = kTrkMaxCtas / kTrkWarps;

enum { TRK_RAYS = 0, TRK_CALIB = 1 };

struct TrackArgs {
  const float* Xf;     // (n,3) frame points, already gathered by the match index (tracker.py:154)
  const float* Xk;     // rays: (n,3) keyframe points; calib: (n,3) measurement [u, v, log z] (tracker.py:147)
  const float* Qk;     // (n)
  const uint8_t* valid;       // (n) bool
  const uint8_t* valid_meas;  // calib only: (n) bool
  const float* K;             // calib only: 3x3 row-major
  const float* T_WCf;         // 8
  const float* T_WCk;         // 8
  long long n;
  int cap;             // shared-memory capacity per CTA, points
  float w0, w1;        // 1/sigma_ray, 1/sigma_dist  resp.  1/sigma_pixel, 1/sigma_depth
  float huber_k;
  int max_iters;
  double rel_error;
  float delta_norm;
  float u_lo, u_hi, v_lo, v_hi, z_eps;
  double* partial;     // [2][grid][kTrkVals], zero at launch (tagged values, see the kernel)
  float* T_WCf_out;
  float* T_CkCf_out;
  int* info;           // [iterations executed, status (0 ok, 1 Cholesky failed)]
  double* cost_out;
};

__device__ __forceinline__ long long trk_ld_relaxed(const long long* p) {
  long long v;
  asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void trk_st_relaxed(long long* p, long long v) {
  asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// dev diagnostics: -DM3R_TRK_TIMELINE prints CTA 0's clock at the stages of the first iterations
#ifdef M3R_TRK_TIMELINE
#define TRK_MARK(k)                                                          \
  do {                                                                       \
    if (blockIdx.x == 0 && tid == 0 && it < 8) s_tl[it][k] = clock64();      \
  } while (0)
#else
#define TRK_MARK(k) \
  do {              \
  } while (0)
#endif

// lietorch Sim3 inverse / product on the [t, q_xyzw, s] layout (tracker.py:180, :212)
__device__ __forceinline__ void sim3_inv(const float* T, float* out) {
  const float qi[4] = {-T[3], -T[4], -T[5], T[6]};
  const float si = 1.0f / T[7];
  float rt[3];
  act_so3_d(qi, T, rt);
  out[0] = -si * rt[0]; out[1] = -si * rt[1]; out[2] = -si * rt[2];
  out[3] = qi[0]; out[4] = qi[1]; out[5] = qi[2]; out[6] = qi[3];
  out[7] = si;
}
__device__ __forceinline__ void sim3_mul(const float* A, const float* B, float* out) {
  float rb[3];
  act_so3_d(A + 3, B, rb);
  out[0] = A[7] * rb[0] + A[0]; out[1] = A[7] * rb[1] + A[1]; out[2] = A[7] * rb[2] + A[2];
  quat_comp_d(A + 3, B + 3, out + 3);
  out[7] = A[7] * B[7];
}

// one weighted residual row: rho2 = sqrt_info^2 * huber(sqrt_info * e)  (tracker.py:157-163)
template <int MASK>
__device__ __forceinline__ void trk_row(float (&acc)[kTrkVals], const float (&J)[7], float sqrt_info, float e, float k) {
  const float wr = fabsf(sqrt_info * e);
  const float hub = wr < k ? 1.0f : __fdividef(k, wr);
  const float rho2 = sqrt_info * sqrt_info * hub;
  float wJ[7];
#pragma unroll
  for (int a = 0; a < 7; a++)
    if ((MASK >> a) & 1) wJ[a] = rho2 * J[a];
  const float we = rho2 * e;
#pragma unroll
  for (int a = 0; a < 7; a++) {
    if (!((MASK >> a) & 1)) continue;
#pragma unroll
    for (int b = 0; b <= a; b++) {
      if (!((MASK >> b) & 1)) continue;
      acc[a * (a + 1) / 2 + b] = fmaf(wJ[a], J[b], acc[a * (a + 1) / 2 + b]);
    }
    acc[28 + a] = fmaf(we, J[a], acc[28 + a]);
  }
  acc[35] = fmaf(we, e, acc[35]);
}

struct TrkPose {
  float t[3], q[4], s;
};

// p = T.act(Xf) (geometry.py:46), r = z - h(x), J = -dh/dp [I | -[p]x | p] (tracker.py:188-193) with the structural
// zeros of the product written out:  N p = 0 and N [p]x = [r]x for N = (I - r r^T)/d, r^T [p]x = 0.
__device__ __forceinline__ void trk_point_rays(float (&acc)[kTrkVals], const TrkPose& T, const float4 a, const float4 b,
                                               float w0, float w1, float k) {
  const float X[3] = {a.x, a.y, a.z};
  float p[3];
  act_so3_d(T.q, X, p);
  p[0] = T.s * p[0] + T.t[0];
  p[1] = T.s * p[1] + T.t[1];
  p[2] = T.s * p[2] + T.t[2];
  const float d2 = p[0] * p[0] + p[1] * p[1] + p[2] * p[2];
  const float dinv = rsqrtf(d2);  // MUFU (2 ulp): far inside the 1e-4 pose tolerance
  const float d = d2 * dinv;
  const float r0 = dinv * p[0], r1 = dinv * p[1], r2 = dinv * p[2];
  const float si_ray = w0 * a.w, si_dist = w1 * a.w;
  {
    const float J[7] = {-dinv * (1.0f - r0 * r0), dinv * r0 * r1, dinv * r0 * r2, 0.0f, -r2, r1, 0.0f};
    trk_row<0x37>(acc, J, si_ray, b.x - r0, k);
  }
  {
    const float J[7] = {dinv * r1 * r0, -dinv * (1.0f - r1 * r1), dinv * r1 * r2, r2, 0.0f, -r0, 0.0f};
    trk_row<0x2F>(acc, J, si_ray, b.y - r1, k);
  }
  {
    const float J[7] = {dinv * r2 * r0, dinv * r2 * r1, -dinv * (1.0f - r2 * r2), -r1, r0, 0.0f, 0.0f};
    trk_row<0x1F>(acc, J, si_ray, b.z - r2, k);
  }
  {
    const float J[7] = {-r0, -r1, -r2, 0.0f, 0.0f, 0.0f, -d};
    trk_row<0x47>(acc, J, si_dist, b.w - d, k);
  }
}

struct TrkCalib {
  float fx, fy, cx, cy, u_lo, u_hi, v_lo, v_hi, z_eps;
};

// project_calib (geometry.py:63-108): pz = [u, v, log z]; J = -dpz/dP [I | -[p]x | p], row a -> [-a, a x p, -a.p]
__device__ __forceinline__ void trk_point_calib(float (&acc)[kTrkVals], const TrkPose& T, const TrkCalib& C,
                                                const float4 a, const float4 b, float w0, float w1, float k) {
  const float X[3] = {a.x, a.y, a.z};
  float p[3];
  act_so3_d(T.q, X, p);
  const float x = T.s * p[0] + T.t[0], y = T.s * p[1] + T.t[1], z = T.s * p[2] + T.t[2];
  const float u = (C.fx * x + C.cx * z) / z, v = (C.fy * y + C.cy * z) / z;
  const bool ok = (u > C.u_lo) && (u < C.u_hi) && (v > C.v_lo) && (v < C.v_hi) && (z > C.z_eps);
  if (!ok) return;  // valid2 = valid_proj & valid_meas_k gates sqrt_info (tracker.py:243-244)
  const float zi = __fdividef(1.0f, z);
  const float xz = x * zi, yz = y * zi;
  const float si_px = w0 * a.w, si_dep = w1 * a.w;
  {
    const float fz = C.fx * zi;
    const float J[7] = {-fz, 0.0f, fz * xz, C.fx * xz * yz, -C.fx * (1.0f + xz * xz), C.fx * yz, 0.0f};
    trk_row<0x3D>(acc, J, si_px, b.x - u, k);
  }
  {
    const float fz = C.fy * zi;
    const float J[7] = {0.0f, -fz, fz * yz, C.fy * (1.0f + yz * yz), -C.fy * xz * yz, -C.fy * xz, 0.0f};
    trk_row<0x3E>(acc, J, si_px, b.y - v, k);
  }
  {
    const float J[7] = {0.0f, 0.0f, -zi, -yz, xz, 0.0f, -1.0f};
    trk_row<0x5C>(acc, J, si_dep, b.z - logf(z), k);
  }
}

__device__ __forceinline__ double trk_rsqrt(double x) {  // fp32 seed + 2 Newton steps in fp64 (as in gn_small.cuh)
  double y = (double)rsqrtf((float)x);
  const double hx = 0.5 * x;
  y = y * (1.5 - hx * y * y);  // 4e-7 -> 2e-13
  y = y * (1.5 - hx * y * y);  // -> below 1 ulp
  return y;
}

// H tau = g for the 7x7 system, H = lower triangle in v[0..27], g = -v[28..34]; returns false if H is not positive
// definite (torch.linalg.cholesky raises there, tracker.py:168).  Right-looking LL^T on the augmented matrix [H | g]
// (row 7 = rhs, so the forward substitution comes for free), every index a compile-time constant: all in registers.
template <int J>
__device__ __forceinline__ void trk_chol_step(double (&L)[36], double (&inv)[7], bool& ok) {
  const double d = L[J * (J + 1) / 2 + J];
  if (!(d > 0.0) || !isfinite(d)) ok = false;
  inv[J] = trk_rsqrt(d);
#pragma unroll
  for (int i = J + 1; i < 8; i++) L[i * (i + 1) / 2 + J] *= inv[J];
#pragma unroll
  for (int i = J + 1; i < 8; i++)
#pragma unroll
    for (int c = J + 1; c <= i && c < 7; c++) L[i * (i + 1) / 2 + c] -= L[i * (i + 1) / 2 + J] * L[c * (c + 1) / 2 + J];
}

__device__ __forceinline__ bool trk_solve7(const double* v, float* tau) {
  double L[36], inv[7];  // rows 0..6 = H, row 7 (L[28..34]) = rhs
#pragma unroll
  for (int i = 0; i < 28; i++) L[i] = v[i];
#pragma unroll
  for (int i = 0; i < 7; i++) L[28 + i] = -v[28 + i];
  L[35] = 0.0;
  bool ok = true;
  trk_chol_step<0>(L, inv, ok);
  trk_chol_step<1>(L, inv, ok);
  trk_chol_step<2>(L, inv, ok);
  trk_chol_step<3>(L, inv, ok);
  trk_chol_step<4>(L, inv, ok);
  trk_chol_step<5>(L, inv, ok);
  trk_chol_step<6>(L, inv, ok);
  // back substitution L^T x = y, y = L[28..34]
  double x[7];
#pragma unroll
  for (int j = 6; j >= 0; j--) {
    double s = L[28 + j];
#pragma unroll
    for (int k = 6; k > j; k--) s -= L[k * (k + 1) / 2 + j] * x[k];
    x[j] = s * inv[j];
  }
#pragma unroll
  for (int j = 0; j < 7; j++) {
    tau[j] = (float)x[j];
    if (!isfinite(x[j])) ok = false;
  }
  return ok;
}

// ============================================================================================== the kernel
template <int KIND>
__global__ void __launch_bounds__(kTrkThreads, 1) track_kernel(const TrackArgs A) {
  extern __shared__ __align__(16) float4 s_pts[];
  float4* sa = s_pts;          // (Xf.xyz, sqrt(Q))
  float4* sb = s_pts + A.cap;  // rays: (ray_k.xyz, dist_k); calib: (u, v, log z, -)
  __shared__ float s_red[kTrkWarps][kTrkVals];
  __shared__ double s_part[kTrkWarps][kTrkVals];
  __shared__ double s_sum[kTrkVals];
  __shared__ float s_T[8];
  __shared__ int s_cnt[kTrkWarps];
  __shared__ int s_stop;
#ifdef M3R_TRK_TIMELINE
  __shared__ long long s_tl[8][8];
#endif
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int G = gridDim.x;

  // ---- load: gate, pre-compute the keyframe side, compact into shared memory in a fixed order.  128-point tiles are
  //      dealt round-robin to the CTAs (tile t -> CTA t mod G), so a CTA's share differs from the mean by < 1 tile
  int M = 0;
  const long long ntiles = (A.n + kTrkTile - 1) / kTrkTile;
  const long long rounds = (ntiles + (long long)G * kTrkTilesPerRound - 1) / ((long long)G * kTrkTilesPerRound);
  for (long long r = 0; r < rounds; r++) {
    const long long tile = (r * kTrkTilesPerRound + tid / kTrkTile) * G + blockIdx.x;
    const long long i = tile * kTrkTile + (tid % kTrkTile);
    float w = 0.0f;
    if (tile < ntiles && i < A.n && A.valid[i]) {
      if (KIND == TRK_RAYS || A.valid_meas[i]) w = sqrtf(__ldg(A.Qk + i));
    }
    const bool keep = !(w == 0.0f);  // a NaN weight is kept and fails the factorisation, as in the reference
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    if (keep) {
      a = make_float4(__ldg(A.Xf + 3 * i), __ldg(A.Xf + 3 * i + 1), __ldg(A.Xf + 3 * i + 2), w);
      const float k0 = __ldg(A.Xk + 3 * i), k1 = __ldg(A.Xk + 3 * i + 1), k2 = __ldg(A.Xk + 3 * i + 2);
      if (KIND == TRK_RAYS) {  // rd_k = point_to_ray_dist(Xk) (tracker.py:183)
        const float d = sqrtf(k0 * k0 + k1 * k1 + k2 * k2);
        const float dinv = 1.0f / d;
        b = make_float4(dinv * k0, dinv * k1, dinv * k2, d);
      } else {
        b = make_float4(k0, k1, k2, 0.0f);
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_cnt[warp] = __popc(bal);
    __syncthreads();
    int off = M, tot = 0;
#pragma unroll
    for (int wv = 0; wv < kTrkWarps; wv++) {
      const int c = s_cnt[wv];
      if (wv < warp) off += c;
      tot += c;
    }
    if (keep) {
      const int o = off + __popc(bal & ((1u << lane) - 1u));
      sa[o] = a;
      sb[o] = b;
    }
    M += tot;
    __syncthreads();
  }

  // ---- T_CkCf = T_WCk^-1 * T_WCf (tracker.py:180)
  if (tid == 0) {
    float Tf[8], Tk[8], Tki[8], Tr[8];
    for (int i = 0; i < 8; i++) {
      Tf[i] = A.T_WCf[i];
      Tk[i] = A.T_WCk[i];
    }
    sim3_inv(Tk, Tki);
    sim3_mul(Tki, Tf, Tr);
    for (int i = 0; i < 8; i++) s_T[i] = Tr[i];
    s_stop = 0;
  }
  TrkCalib C;
  if (KIND == TRK_CALIB) {
    C.fx = __ldg(A.K + 0); C.cx = __ldg(A.K + 2); C.fy = __ldg(A.K + 4); C.cy = __ldg(A.K + 5);
    C.u_lo = A.u_lo; C.u_hi = A.u_hi; C.v_lo = A.v_lo; C.v_hi = A.v_hi; C.z_eps = A.z_eps;
  }
  __syncthreads();

  double old_cost = __longlong_as_double(0x7ff0000000000000LL);  // +inf (tracker.py:185)
  int it = 0, status = 0;
  double new_cost = 0.0;
  for (; it < A.max_iters; it++) {
    TrkPose T;
    T.t[0] = s_T[0]; T.t[1] = s_T[1]; T.t[2] = s_T[2];
    T.q[0] = s_T[3]; T.q[1] = s_T[4]; T.q[2] = s_T[5]; T.q[3] = s_T[6];
    T.s = s_T[7];
    float acc[kTrkVals];
#pragma unroll
    for (int i = 0; i < kTrkVals; i++) acc[i] = 0.0f;
    for (int i = tid; i < M; i += kTrkThreads) {
      if (KIND == TRK_RAYS)
        trk_point_rays(acc, T, sa[i], sb[i], A.w0, A.w1, A.huber_k);
      else
        trk_point_calib(acc, T, C, sa[i], sb[i], A.w0, A.w1, A.huber_k);
    }
    TRK_MARK(0);
    // lanes: transposing butterfly for acc[0..31] (lane l ends with the warp total of acc[l]), xor tree for the rest
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
      const bool upper = (lane & s) != 0;
#pragma unroll
      for (int i = 0; i < s; i++) {
        const float send = upper ? acc[i] : acc[i + s];
        const float keep = upper ? acc[i + s] : acc[i];
        acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
      }
    }
    s_red[warp][lane] = acc[0];
#pragma unroll
    for (int i = 32; i < kTrkVals; i++) {
      float v = acc[i];
      v += __shfl_xor_sync(0xffffffffu, v, 16);
      v += __shfl_xor_sync(0xffffffffu, v, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 4);
      v += __shfl_xor_sync(0xffffffffu, v, 2);
      v += __shfl_xor_sync(0xffffffffu, v, 1);
      if (lane == 0) s_red[warp][i] = v;
    }
    __syncthreads();
    // warps -> CTA partial in fp64, published WITHOUT a flag or a fence: the two low mantissa bits of every double
    // carry the iteration tag (it + 1) mod 4, and readers poll each value until its tag matches.  A slot of buffer
    // it & 1 last held iteration it - 2 (tag differs by 2) or the zero fill (tag 0 != the tags 1, 2 of iterations 0,
    // 1), and it is rewritten for it + 2 only after its owner has gathered it + 1, i.e. after every CTA has finished
    // reading it.  Each 8-byte value validates itself, so no ordering between different addresses is needed.  Cost of
    // the earlier schemes (r1): per-CTA flags polled by everyone 14 us/iteration, arrival counter + release 5.2 us.
    const long long tag = (it + 1) & 3;
    long long* mine = reinterpret_cast<long long*>(A.partial) + ((size_t)(it & 1) * G + blockIdx.x) * kTrkVals;
    if (tid < kTrkVals) {
      double v = 0.0;
#pragma unroll
      for (int wv = 0; wv < kTrkWarps; wv++) v += (double)s_red[wv][tid];
      trk_st_relaxed(mine + tid, (__double_as_longlong(v) & ~3LL) | tag);
    }
    TRK_MARK(1);
    {
      const long long* base = reinterpret_cast<const long long*>(A.partial) + (size_t)(it & 1) * G * kTrkVals;
      long long b0[kTrkPerWarp], b1[kTrkPerWarp];
      bool ok;
      do {
        ok = true;
#pragma unroll
        for (int j = 0; j < kTrkPerWarp; j++) {
          const int c = warp + kTrkWarps * j;
          b0[j] = tag;
          b1[j] = tag;
          if (c < G) {
            b0[j] = trk_ld_relaxed(base + (size_t)c * kTrkVals + lane);
            if (lane < kTrkVals - 32) b1[j] = trk_ld_relaxed(base + (size_t)c * kTrkVals + 32 + lane);
          }
        }
#pragma unroll
        for (int j = 0; j < kTrkPerWarp; j++) ok = ok && ((b0[j] & 3) == tag) && ((b1[j] & 3) == tag);
      } while (!__all_sync(0xffffffffu, ok));
      TRK_MARK(2);
      double v0[kTrkPerWarp], v1[kTrkPerWarp];
#pragma unroll
      for (int j = 0; j < kTrkPerWarp; j++) {
        const int c = warp + kTrkWarps * j;
        v0[j] = c < G ? __longlong_as_double(b0[j]) : 0.0;
        v1[j] = (c < G && lane < kTrkVals - 32) ? __longlong_as_double(b1[j]) : 0.0;
      }
      double a0 = 0.0, a1 = 0.0;
#pragma unroll
      for (int j = 0; j < kTrkPerWarp; j++) {
        a0 += v0[j];
        a1 += v1[j];
      }
      s_part[warp][lane] = a0;
      if (lane < kTrkVals - 32) s_part[warp][32 + lane] = a1;
    }
    __syncthreads();
    if (tid < kTrkVals) {
      double v = 0.0;
#pragma unroll
      for (int wv = 0; wv < kTrkWarps; wv++) v += s_part[wv][tid];
      s_sum[tid] = v;
    }
    __syncthreads();
    TRK_MARK(3);
    // ---- solve, retract, convergence test (redundantly, identically, on every CTA)
    if (tid == 0) {
      float tau[7];
      const bool ok = trk_solve7(s_sum, tau);
      if (!ok) {
        s_stop = 2;
      } else {
        float T8[8];
        for (int i = 0; i < 8; i++) T8[i] = s_T[i];
        retr_sim3(tau, T8);  // T_CkCf.retr(tau) (tracker.py:196)
        for (int i = 0; i < 8; i++) s_T[i] = T8[i];
        // check_convergence (nonlinear_optimizer.py:5-25): cost = 0.5 b^T b (tracker.py:166)
        const double cost = 0.5 * s_sum[35];
        const double rel_dec = fabs((old_cost - cost) / old_cost);  // NaN on the first iteration -> false
        float nrm = 0.0f;
        for (int i = 0; i < 7; i++) nrm = fmaf(tau[i], tau[i], nrm);
        nrm = sqrtf(nrm);
        if (rel_dec < A.rel_error || nrm < A.delta_norm) s_stop = 1;
      }
    }
    TRK_MARK(4);
    __syncthreads();
    new_cost = 0.5 * s_sum[35];
    const int stop = s_stop;
    if (stop == 2) {
      status = 1;
      break;
    }
    old_cost = new_cost;
    if (stop == 1) {
      it++;
      break;
    }
  }
  // ---- T_WCf = T_WCk * T_CkCf (tracker.py:212)
  if (blockIdx.x == 0 && tid == 0) {
    float Tk[8], Tr[8], Tw[8];
    for (int i = 0; i < 8; i++) {
      Tk[i] = A.T_WCk[i];
      Tr[i] = s_T[i];
    }
    sim3_mul(Tk, Tr, Tw);
    for (int i = 0; i < 8; i++) {
      A.T_WCf_out[i] = Tw[i];
      A.T_CkCf_out[i] = Tr[i];
    }
#ifdef M3R_TRK_TIMELINE
    for (int k = 0; k < 8 && k < it; k++)
      printf("it %d: compute %lld reduce+publish %lld wait %lld gather %lld solve %lld | next %lld\n", k,
             s_tl[k][0] - (k ? s_tl[k - 1][4] : s_tl[k][0]), s_tl[k][1] - s_tl[k][0], s_tl[k][2] - s_tl[k][1],
             s_tl[k][3] - s_tl[k][2], s_tl[k][4] - s_tl[k][3], k + 1 < it && k < 7 ? s_tl[k + 1][0] - s_tl[k][4] : 0LL);
#endif
    A.info[0] = it;
    A.info[1] = status;
    A.cost_out[0] = new_cost;
  }
}

// ============================================================================================== launcher
static int track_launch(int kind, TrackArgs A, cudaStream_t st) {
  M3R_REQUIRE(A.n >= 0, "track: negative point count");
  M3R_REQUIRE(A.max_iters >= 0, "track: max_iters must be >= 0");
  M3R_REQUIRE(A.T_WCf && A.T_WCk && A.T_WCf_out && A.T_CkCf_out && A.info && A.cost_out, "track: null pointer");
  M3R_REQUIRE(A.n == 0 || (A.Xf && A.Xk && A.Qk && A.valid), "track: null pointer");
  if (kind == TRK_CALIB) M3R_REQUIRE(A.K && (A.n == 0 || A.valid_meas), "track: null pointer (calib)");
  const long long ntiles = ceil_div_ll(A.n, kTrkTile);
  int G = sm_count();
  if (G > kTrkMaxCtas) G = kTrkMaxCtas;
  if (ceil_div_ll(ntiles, kTrkTilesPerRound) < G) G = ntiles > 0 ? (int)ceil_div_ll(ntiles, kTrkTilesPerRound) : 1;
  const long long rounds = ntiles > 0 ? ceil_div_ll(ntiles, (long long)G * kTrkTilesPerRound) : 1;
  const long long cap = rounds * kTrkThreads;
  const size_t smem = (size_t)cap * 2 * sizeof(float4);
  static int smem_max_dev[kMaxDevices] = {};  // per device: the opt-in applies to the current device only
  int& smem_max = smem_max_dev[current_device_slot()];
  if (smem_max == 0) {
    int dev = 0, v = 0;
    M3R_CUDA(cudaGetDevice(&dev));
    M3R_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    v -= 8 * 1024;  // static shared memory of the kernel + reserve
    M3R_CUDA(cudaFuncSetAttribute(track_kernel<TRK_RAYS>, cudaFuncAttributeMaxDynamicSharedMemorySize, v));
    M3R_CUDA(cudaFuncSetAttribute(track_kernel<TRK_CALIB>, cudaFuncAttributeMaxDynamicSharedMemorySize, v));
    smem_max = v;
  }
  M3R_REQUIRE(smem <= (size_t)smem_max,
              "track: %lld points exceed the shared-memory residency of this device (%lld points)", A.n,
              (long long)(smem_max / 32) * G);
  A.cap = (int)cap;
  {
    const int rc = keep_async_pool();
    if (rc != M3R_OK) return rc;
  }
  const size_t pbytes = align_up((size_t)2 * G * kTrkVals * sizeof(double), 256);
  void* scratch = nullptr;
  M3R_CUDA(cudaMallocAsync(&scratch, pbytes, st));
  A.partial = (double*)scratch;
  M3R_CUDA(cudaMemsetAsync(scratch, 0, pbytes, st));
  void* params[] = {(void*)&A};
  const void* fn = kind == TRK_RAYS ? (const void*)track_kernel<TRK_RAYS> : (const void*)track_kernel<TRK_CALIB>;
  // cooperative launch: the CTAs spin on each other's partial sums, so they must be co-resident
  M3R_CUDA(cudaLaunchCooperativeKernel(fn, dim3(G), dim3(kTrkThreads), params, smem, st));
  count_launch();
  M3R_CUDA(cudaFreeAsync(scratch, st));
  return M3R_OK;
}

}  // namespace m3r

using namespace m3r;

extern "C" int m3r_track_rays(const float* Xf, const float* Xk, const float* T_WCf, const float* T_WCk,
                              const float* Qk, const uint8_t* valid, long long n, double sigma_ray, double sigma_dist,
                              double huber_k, int max_iters, double rel_error, double delta_norm, float* T_WCf_out,
                              float* T_CkCf_out, int* info, double* cost_out, void* stream) {
  M3R_REQUIRE(sigma_ray > 0 && sigma_dist > 0, "track_rays: sigmas must be positive");
  TrackArgs A = {};
  A.Xf = Xf; A.Xk = Xk; A.Qk = Qk; A.valid = valid;
  A.T_WCf = T_WCf; A.T_WCk = T_WCk;
  A.n = n;
  A.w0 = (float)(1.0 / sigma_ray);  // `1 / sigma * valid` is a float32 tensor (tracker.py:175-176)
  A.w1 = (float)(1.0 / sigma_dist);
  A.huber_k = (float)huber_k;
  A.max_iters = max_iters;
  A.rel_error = rel_error;
  A.delta_norm = (float)delta_norm;
  A.T_WCf_out = T_WCf_out; A.T_CkCf_out = T_CkCf_out; A.info = info; A.cost_out = cost_out;
  return track_launch(TRK_RAYS, A, (cudaStream_t)stream);
}

extern "C" int m3r_track_calib(const float* Xf, const float* meas_k, const float* T_WCf, const float* T_WCk,
                               const float* Qk, const uint8_t* valid, const uint8_t* valid_meas_k, const float* K,
                               long long n, int img_h, int img_w, double sigma_pixel, double sigma_depth,
                               double pixel_border, double depth_eps, double huber_k, int max_iters, double rel_error,
                               double delta_norm, float* T_WCf_out, float* T_CkCf_out, int* info, double* cost_out,
                               void* stream) {
  M3R_REQUIRE(sigma_pixel > 0 && sigma_depth > 0, "track_calib: sigmas must be positive");
  TrackArgs A = {};
  A.Xf = Xf; A.Xk = meas_k; A.Qk = Qk; A.valid = valid; A.valid_meas = valid_meas_k; A.K = K;
  A.T_WCf = T_WCf; A.T_WCk = T_WCk;
  A.n = n;
  A.w0 = (float)(1.0 / sigma_pixel);
  A.w1 = (float)(1.0 / sigma_depth);
  A.huber_k = (float)huber_k;
  A.max_iters = max_iters;
  A.rel_error = rel_error;
  A.delta_norm = (float)delta_norm;
  // valid_u = (u > border) & (u < img_w - 1 - border) (geometry.py:77-78); python scalars compare against float32 u
  A.u_lo = (float)pixel_border; A.u_hi = (float)((double)img_w - 1.0 - pixel_border);
  A.v_lo = (float)pixel_border; A.v_hi = (float)((double)img_h - 1.0 - pixel_border);
  A.z_eps = (float)depth_eps;
  A.T_WCf_out = T_WCf_out; A.T_CkCf_out = T_CkCf_out; A.info = info; A.cost_out = cost_out;
  return track_launch(TRK_CALIB, A, (cudaStream_t)stream);
}
