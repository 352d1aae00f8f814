// Batched action consumer for arm bodies: what happens to policy_action right after infer_policy.
//   RolloutBase.set_command_data           robo_manip_baselines/common/base/RolloutBase.py:496-509
//   ArmManager.set_command_data & friends  robo_manip_baselines/common/body/ArmManager.py:88-160
//   ArmManager.inverse_kinematics          robo_manip_baselines/common/body/ArmManager.py:230-254
// One thread = one episode, fp64 throughout (the reference is numpy / Pinocchio in double). Per episode the state is the
// arm joint positions, the gripper joint positions and the target SE3. The kinematic model is a serial chain of
// revolute joints given by value (joint placements + axes, what pin.buildModelFromUrdf yields for the arm URDFs of the
// reference). Pinocchio's algorithms are restated (forward kinematics, LOCAL joint Jacobian, log6 / Jlog6 of
// explog.hpp, integrate = q + dq for revolute joints); oracle/ik_oracle.py is the numpy twin the tests compare with.
// This is latency-trivial work (a 6 x 6 solve per episode); it lives on the device so that the (B, Da) action matrix
// never has to visit the host between the policy and a device-side consumer.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/rmb_b200.h"

namespace rmb {

struct Se3 { double R[9]; double p[3]; };   // R row-major

__device__ __forceinline__ Se3 se3_mul(const Se3& a, const Se3& b) {
  Se3 o;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j) o.R[3 * i + j] = a.R[3 * i] * b.R[j] + a.R[3 * i + 1] * b.R[3 + j] + a.R[3 * i + 2] * b.R[6 + j];
    o.p[i] = a.R[3 * i] * b.p[0] + a.R[3 * i + 1] * b.p[1] + a.R[3 * i + 2] * b.p[2] + a.p[i];
  }
  return o;
}
__device__ __forceinline__ Se3 se3_inv(const Se3& a) {
  Se3 o;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j) o.R[3 * i + j] = a.R[3 * j + i];
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) o.p[i] = -(o.R[3 * i] * a.p[0] + o.R[3 * i + 1] * a.p[1] + o.R[3 * i + 2] * a.p[2]);
  return o;
}
__device__ __forceinline__ void cross3(const double* a, const double* b, double* o) {
  o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}

constexpr double kIkPrec = 0.01;   // 1e-8 ** 0.25, the Taylor switch of oracle/ik_oracle.py (PREC)

// pinocchio log3: rotation matrix -> axis * angle, angle
__device__ inline void log3_dev(const double* R, double* w, double* theta_out) {
  const double tr = R[0] + R[4] + R[8];
  double theta;
  if (tr >= 3.0) theta = 0.0;
  else if (tr <= -1.0) theta = 3.14159265358979323846;
  else theta = acos((tr - 1.0) / 2.0);
  if (theta >= 3.14159265358979323846 - 1e-2) {
    const double cphi = -(tr - 1.0) / 2.0;
    const double beta = theta * theta / (1.0 + cphi);
    const double d[3] = {R[0], R[4], R[8]};
    const double s[3] = {R[7] > R[5] ? 1.0 : -1.0, R[2] > R[6] ? 1.0 : -1.0, R[3] > R[1] ? 1.0 : -1.0};
#pragma unroll
    for (int i = 0; i < 3; ++i) w[i] = s[i] * sqrt(fmax((d[i] + cphi) * beta, 0.0));
  } else {
    const double t = theta > kIkPrec ? theta / sin(theta) : 1.0 + theta * theta / 6.0;
    w[0] = 0.5 * t * (R[7] - R[5]); w[1] = 0.5 * t * (R[2] - R[6]); w[2] = 0.5 * t * (R[3] - R[1]);
  }
  *theta_out = theta;
}

// pinocchio log6: SE3 -> [v; w]
__device__ inline void log6_dev(const Se3& M, double* out) {
  double w[3], t;
  log3_dev(M.R, w, &t);
  const double t2 = t * t;
  double alpha, beta;
  if (t < kIkPrec) { alpha = 1.0 - t2 / 12.0 - t2 * t2 / 720.0; beta = 1.0 / 12.0 + t2 / 720.0; }
  else { const double st = sin(t), ct = cos(t); alpha = t * st / (2.0 * (1.0 - ct)); beta = 1.0 / t2 - st / (2.0 * t * (1.0 - ct)); }
  double wxp[3];
  cross3(w, M.p, wxp);
  const double wp = w[0] * M.p[0] + w[1] * M.p[1] + w[2] * M.p[2];
#pragma unroll
  for (int i = 0; i < 3; ++i) { out[i] = alpha * M.p[i] - 0.5 * wxp[i] + beta * wp * w[i]; out[3 + i] = w[i]; }
}

// pinocchio Jlog6 (6 x 6, row-major)
__device__ inline void jlog6_dev(const Se3& M, double* J) {
  double w[3], t;
  log3_dev(M.R, w, &t);
  double A[9];
  {
    double alpha, diag;
    if (t < kIkPrec) { alpha = 1.0 / 12.0 + t * t / 720.0; diag = 0.5 * (2.0 - t * t / 6.0); }
    else { const double st = sin(t), ct = cos(t), s1 = st / (1.0 - ct); alpha = 1.0 / (t * t) - s1 / (2.0 * t); diag = 0.5 * (t * s1); }
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) A[3 * i + j] = alpha * w[i] * w[j] + (i == j ? diag : 0.0);
    A[1] += -0.5 * w[2]; A[2] += 0.5 * w[1]; A[3] += 0.5 * w[2]; A[5] += -0.5 * w[0]; A[6] += -0.5 * w[1]; A[7] += 0.5 * w[0];
  }
  const double t2 = t * t;
  double beta, bdot;
  if (t < kIkPrec) { beta = 1.0 / 12.0 + t2 / 720.0; bdot = 1.0 / 360.0; }
  else {
    const double tinv = 1.0 / t, t2inv = tinv * tinv, st = sin(t), ct = cos(t), inv = 1.0 / (2.0 * (1.0 - ct));
    beta = t2inv - st * tinv * inv;
    bdot = -2.0 * t2inv * t2inv + (1.0 + st * tinv) * t2inv * inv;
  }
  const double* p = M.p;
  const double wp = w[0] * p[0] + w[1] * p[1] + w[2] * p[2];
  double v3[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) v3[i] = (bdot * wp) * w[i] - (t2 * bdot + 2.0 * beta) * p[i];
  double C[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C[3 * i + j] = v3[i] * w[j] + beta * w[i] * p[j] + (i == j ? wp * beta : 0.0);
  C[1] += -0.5 * p[2]; C[2] += 0.5 * p[1]; C[3] += 0.5 * p[2]; C[5] += -0.5 * p[0]; C[6] += -0.5 * p[1]; C[7] += 0.5 * p[0];
#pragma unroll
  for (int i = 0; i < 36; ++i) J[i] = 0.0;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      J[6 * i + j] = A[3 * i + j];
      J[6 * (3 + i) + 3 + j] = A[3 * i + j];
      J[6 * i + 3 + j] = C[3 * i] * A[j] + C[3 * i + 1] * A[3 + j] + C[3 * i + 2] * A[6 + j];
    }
}

// forward kinematics of joints 1..n (0-based array); returns the frame of joint `jid` (1-based)
__device__ inline Se3 arm_fk(const rmb_arm_model_t& m, const double* q, Se3* frames) {
  Se3 cur;
#pragma unroll
  for (int i = 0; i < 9; ++i) cur.R[i] = (i % 4 == 0) ? 1.0 : 0.0;
  cur.p[0] = cur.p[1] = cur.p[2] = 0.0;
  for (int k = 0; k < m.n_joints; ++k) {
    Se3 pl;
#pragma unroll
    for (int i = 0; i < 9; ++i) pl.R[i] = m.placement_R[k][i];
#pragma unroll
    for (int i = 0; i < 3; ++i) pl.p[i] = m.placement_p[k][i];
    const double* a = m.axis[k];
    const double s = sin(q[k]), c1 = 1.0 - cos(q[k]);
    const double K[9] = {0.0, -a[2], a[1], a[2], 0.0, -a[0], -a[1], a[0], 0.0};
    Se3 rq;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const double kk = K[3 * i] * K[j] + K[3 * i + 1] * K[3 + j] + K[3 * i + 2] * K[6 + j];
        rq.R[3 * i + j] = (i == j ? 1.0 : 0.0) + s * K[3 * i + j] + c1 * kk;
      }
    rq.p[0] = rq.p[1] = rq.p[2] = 0.0;
    cur = se3_mul(se3_mul(cur, pl), rq);
    if (frames != nullptr) frames[k] = cur;
  }
  return (frames != nullptr) ? frames[m.ik_joint_id - 1] : cur;
}

// one damped least-squares step (ArmManager.inverse_kinematics): q <- q + dq
__device__ inline void arm_ik_step(const rmb_arm_model_t& m, double* q, const Se3& target) {
  Se3 frames[RMB_ARM_MAX_JOINTS];
  const Se3 cur = arm_fk(m, q, frames);
  const Se3 err = se3_mul(se3_inv(cur), target);
  double e[6];
  log6_dev(err, e);
  const int n = m.n_joints, jid = m.ik_joint_id;
  double Jl[6 * RMB_ARM_MAX_JOINTS];
  const Se3 inv_j = se3_inv(cur);
  for (int k = 0; k < n; ++k) {
    double col[6] = {0, 0, 0, 0, 0, 0};
    if (k < jid) {
      const Se3 rel = se3_mul(inv_j, frames[k]);
      double a[3];
#pragma unroll
      for (int i = 0; i < 3; ++i) a[i] = rel.R[3 * i] * m.axis[k][0] + rel.R[3 * i + 1] * m.axis[k][1] + rel.R[3 * i + 2] * m.axis[k][2];
      cross3(rel.p, a, col);
      col[3] = a[0]; col[4] = a[1]; col[5] = a[2];
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) Jl[i * RMB_ARM_MAX_JOINTS + k] = col[i];
  }
  double JL[36];
  jlog6_dev(se3_inv(err), JL);
  double J[6 * RMB_ARM_MAX_JOINTS];                    // J = -Jlog6(err^-1) * J_local
  for (int i = 0; i < 6; ++i)
    for (int k = 0; k < n; ++k) {
      double acc = 0.0;
#pragma unroll
      for (int l = 0; l < 6; ++l) acc += JL[6 * i + l] * Jl[l * RMB_ARM_MAX_JOINTS + k];
      J[i * RMB_ARM_MAX_JOINTS + k] = -acc;
    }
  double ee = 0.0;
#pragma unroll
  for (int i = 0; i < 6; ++i) ee += e[i] * e[i];
  const double lam = ee + 1e-6;
  double Amat[36], y[6];
  for (int i = 0; i < 6; ++i) {
    for (int j = 0; j < 6; ++j) {
      double acc = 0.0;
      for (int k = 0; k < n; ++k) acc += J[i * RMB_ARM_MAX_JOINTS + k] * J[j * RMB_ARM_MAX_JOINTS + k];
      Amat[6 * i + j] = acc + (i == j ? lam : 0.0);
    }
    y[i] = e[i];
  }
  // solve Amat y = e: Gaussian elimination with partial pivoting (what LAPACK's gesv behind np.linalg.solve does)
  for (int c = 0; c < 6; ++c) {
    int piv = c;
    double best = fabs(Amat[6 * c + c]);
    for (int r = c + 1; r < 6; ++r) if (fabs(Amat[6 * r + c]) > best) { best = fabs(Amat[6 * r + c]); piv = r; }
    if (piv != c) {
      for (int j = 0; j < 6; ++j) { const double t = Amat[6 * c + j]; Amat[6 * c + j] = Amat[6 * piv + j]; Amat[6 * piv + j] = t; }
      const double t = y[c]; y[c] = y[piv]; y[piv] = t;
    }
    const double inv = 1.0 / Amat[6 * c + c];
    for (int r = c + 1; r < 6; ++r) {
      const double f = Amat[6 * r + c] * inv;
      for (int j = c; j < 6; ++j) Amat[6 * r + j] -= f * Amat[6 * c + j];
      y[r] -= f * y[c];
    }
  }
  for (int c = 5; c >= 0; --c) {
    double acc = y[c];
    for (int j = c + 1; j < 6; ++j) acc -= Amat[6 * c + j] * y[j];
    y[c] = acc / Amat[6 * c + c];
  }
  for (int k = 0; k < n; ++k) {
    double acc = 0.0;
#pragma unroll
    for (int i = 0; i < 6; ++i) acc += J[i * RMB_ARM_MAX_JOINTS + k] * y[i];
    q[k] -= acc;                                         // dq = -J^T y ; pin.integrate = q + dq
  }
}

__device__ inline Se3 se3_from_pose_dev(const double* pose) {   // (tx, ty, tz, qw, qx, qy, qz), Eigen's toRotationMatrix
  const double w = pose[3], x = pose[4], y = pose[5], z = pose[6];
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
  Se3 o;
  o.R[0] = 1 - (tyy + tzz); o.R[1] = txy - twz; o.R[2] = txz + twy;
  o.R[3] = txy + twz; o.R[4] = 1 - (txx + tzz); o.R[5] = tyz - twx;
  o.R[6] = txz - twy; o.R[7] = tyz + twx; o.R[8] = 1 - (txx + tyy);
  o.p[0] = pose[0]; o.p[1] = pose[1]; o.p[2] = pose[2];
  return o;
}
__device__ inline Se3 se3_from_rel_pose_dev(const double* rel) {  // (tx, ty, tz, roll, pitch, yaw): Rz(yaw) Ry(pitch) Rx(roll)
  const double cr = cos(rel[3]), sr = sin(rel[3]), cp = cos(rel[4]), sp = sin(rel[4]), cy = cos(rel[5]), sy = sin(rel[5]);
  Se3 o;
  o.R[0] = cy * cp; o.R[1] = cy * sp * sr - sy * cr; o.R[2] = cy * sp * cr + sy * sr;
  o.R[3] = sy * cp; o.R[4] = sy * sp * sr + cy * cr; o.R[5] = sy * sp * cr - cy * sr;
  o.R[6] = -sp; o.R[7] = cp * sr; o.R[8] = cp * cr;
  o.p[0] = rel[0]; o.p[1] = rel[1]; o.p[2] = rel[2];
  return o;
}
__device__ inline void pose_from_se3_dev(const Se3& s, double* pose) {   // Eigen's matrix -> quaternion branches
  const double* R = s.R;
  pose[0] = s.p[0]; pose[1] = s.p[1]; pose[2] = s.p[2];
  double t = R[0] + R[4] + R[8];
  if (t > 0) {
    t = sqrt(t + 1.0);
    pose[3] = 0.5 * t;
    t = 0.5 / t;
    pose[4] = (R[7] - R[5]) * t; pose[5] = (R[2] - R[6]) * t; pose[6] = (R[3] - R[1]) * t;
    return;
  }
  int i = 0;
  if (R[4] > R[0]) i = 1;
  if (R[8] > R[4 * i]) i = 2;
  const int j = (i + 1) % 3, k = (i + 2) % 3;
  t = sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
  pose[4 + i] = 0.5 * t;
  t = 0.5 / t;
  pose[3] = (R[3 * k + j] - R[3 * j + k]) * t;
  pose[4 + j] = (R[3 * j + i] + R[3 * i + j]) * t;
  pose[4 + k] = (R[3 * k + i] + R[3 * i + k]) * t;
}

// key: 0 command_joint_pos, 1 command_joint_pos_rel, 2 command_gripper_joint_pos, 3 command_eef_pose, 4 command_eef_pose_rel,
//      5 reset (arm <- init, gripper <- init, target <- FK(init))
__global__ void arm_command_kernel(rmb_arm_model_t m, rmb_arm_layout_t lay, int key, const double* __restrict__ command,
                                   int cmd_stride, int is_skip, const uint8_t* __restrict__ mask, double* __restrict__ arm_q,
                                   double* __restrict__ grip_q, double* __restrict__ target, int B) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  if (mask != nullptr && mask[b] == 0) return;
  const int n = m.n_joints, ng = lay.n_gripper;
  double q[RMB_ARM_MAX_JOINTS];
  for (int k = 0; k < n; ++k) q[k] = arm_q[b * n + k];
  double* g = grip_q + b * ng;
  Se3 tg;
  for (int i = 0; i < 9; ++i) tg.R[i] = target[b * 12 + i];
  for (int i = 0; i < 3; ++i) tg.p[i] = target[b * 12 + 9 + i];
  const double* cmd = (command != nullptr) ? command + static_cast<long long>(b) * cmd_stride : nullptr;
  auto clip_grip = [&](int i, double v) { g[i] = fmin(fmax(v, lay.gripper_low[i]), lay.gripper_high[i]); };
  if (key == 5) {
    for (int k = 0; k < n; ++k) q[k] = lay.init_arm[k];
    for (int i = 0; i < ng; ++i) g[i] = lay.init_gripper[i];
    tg = arm_fk(m, q, nullptr);
    if (m.ik_joint_id != n) { Se3 fr[RMB_ARM_MAX_JOINTS]; tg = arm_fk(m, q, fr); }
  } else if (key == 0 || key == 1) {
    for (int k = 0; k < n; ++k) {
      const double c = cmd[lay.arm_idx[k]];
      q[k] = (key == 0) ? c : (is_skip ? q[k] : q[k] + c);
    }
    for (int i = 0; i < ng; ++i) {
      const double c = cmd[lay.gripper_idx[i]];
      clip_grip(i, (key == 0) ? c : (is_skip ? g[i] : g[i] + c));
    }
    Se3 fr[RMB_ARM_MAX_JOINTS];
    tg = arm_fk(m, q, fr);                               // target_se3 = current_se3 (ArmManager.py:127-129)
  } else if (key == 2) {
    for (int i = 0; i < ng; ++i) clip_grip(i, cmd[lay.gripper_idx_in_gripper[i]]);
  } else if (key == 3) {
    tg = se3_from_pose_dev(cmd + 7 * lay.eef_idx);
    arm_ik_step(m, q, tg);
  } else if (key == 4) {
    tg = se3_mul(tg, se3_from_rel_pose_dev(cmd + 6 * lay.eef_idx));
    arm_ik_step(m, q, tg);
  }
  for (int k = 0; k < n; ++k) arm_q[b * n + k] = q[k];
  for (int i = 0; i < 9; ++i) target[b * 12 + i] = tg.R[i];
  for (int i = 0; i < 3; ++i) target[b * 12 + 9 + i] = tg.p[i];
}

// get_command_data: COMMAND_JOINT_POS -> (arm, gripper) scattered to their indices; COMMAND_EEF_POSE -> pose of target
__global__ void arm_get_command_kernel(rmb_arm_layout_t lay, int n, const double* __restrict__ arm_q, const double* __restrict__ grip_q,
                                       const double* __restrict__ target, double* __restrict__ joint_pos, int jp_stride,
                                       double* __restrict__ eef_pose, int B) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  if (joint_pos != nullptr) {
    for (int k = 0; k < n; ++k) joint_pos[static_cast<long long>(b) * jp_stride + lay.arm_idx[k]] = arm_q[b * n + k];
    for (int i = 0; i < lay.n_gripper; ++i) joint_pos[static_cast<long long>(b) * jp_stride + lay.gripper_idx[i]] = grip_q[b * lay.n_gripper + i];
  }
  if (eef_pose != nullptr) {
    Se3 tg;
    for (int i = 0; i < 9; ++i) tg.R[i] = target[b * 12 + i];
    for (int i = 0; i < 3; ++i) tg.p[i] = target[b * 12 + 9 + i];
    pose_from_se3_dev(tg, eef_pose + static_cast<long long>(b) * 7);
  }
}

}  // namespace rmb

extern "C" {

int rmb_arm_command(const rmb_arm_model_t* model, const rmb_arm_layout_t* layout, int key, const double* d_command,
                    int cmd_stride, int is_skip, const uint8_t* d_mask, double* d_arm_q, double* d_gripper_q,
                    double* d_target_se3, int batch, void* stream_) {
  if (!model || !layout || !d_arm_q || !d_gripper_q || !d_target_se3) return fail("rmb_arm_command: null argument");
  if (key < 0 || key > 5) return fail("rmb_arm_command: key must be 0..5");
  if (key != 5 && !d_command) return fail("rmb_arm_command: command required");
  if (model->n_joints < 1 || model->n_joints > RMB_ARM_MAX_JOINTS || model->ik_joint_id < 1 || model->ik_joint_id > model->n_joints)
    return fail("rmb_arm_command: bad arm model");
  if (layout->n_gripper < 0 || layout->n_gripper > RMB_ARM_MAX_GRIPPER) return fail("rmb_arm_command: bad layout");
  if (batch < 1) return 0;
  rmb::arm_command_kernel<<<(batch + 63) / 64, 64, 0, static_cast<cudaStream_t>(stream_)>>>(
      *model, *layout, key, d_command, cmd_stride, is_skip, d_mask, d_arm_q, d_gripper_q, d_target_se3, batch);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

int rmb_arm_get_command(const rmb_arm_model_t* model, const rmb_arm_layout_t* layout, const double* d_arm_q,
                        const double* d_gripper_q, const double* d_target_se3, double* d_joint_pos, int joint_pos_stride,
                        double* d_eef_pose, int batch, void* stream_) {
  if (!model || !layout || !d_arm_q || !d_gripper_q || !d_target_se3) return fail("rmb_arm_get_command: null argument");
  if (batch < 1) return 0;
  rmb::arm_get_command_kernel<<<(batch + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream_)>>>(
      *layout, model->n_joints, d_arm_q, d_gripper_q, d_target_se3, d_joint_pos, joint_pos_stride, d_eef_pose, batch);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
