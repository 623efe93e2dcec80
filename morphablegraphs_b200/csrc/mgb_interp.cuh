// The constraint-program interpreter shared by the SIMT and the tensor-core score kernels:
// a depth-first walk over the needed part of the skeleton (forward kinematics in quaternion
// form) with the constraint terms attached to the nodes they read (op list: mgb_internal.cuh).
//   quad(q, out[CPT])      -> the 4 pose channels of contracted quad q for each candidate
//   emit(c, term, w, e)    -> term `term` of candidate c evaluated to e with weight w
// What it replaces per candidate: MotionSpline.evaluate + SkeletonNode.get_global_matrix +
// constraint.evaluate_motion_spline (reference motion_primitive_constraints.py:118-122 and the
// keyframe constraint classes cited in include/mgb200.h).
#pragma once
#include <math.h>

#include "mgb_device.cuh"
#include "mgb_internal.cuh"

__device__ __forceinline__ float masked_dist(unsigned mask, const float* tgt, float px, float py, float pz) {
    float d = 0.f;
    if (mask & 1u) { float e = tgt[0] - px; d = fmaf(e, e, d); }
    if (mask & 2u) { float e = tgt[1] - py; d = fmaf(e, e, d); }
    if (mask & 4u) { float e = tgt[2] - pz; d = fmaf(e, e, d); }
    return sqrtf(d);
}

__device__ __forceinline__ float dist3(const float* a, const float* b) {
    float x = a[0] - b[0], y = a[1] - b[1], z = a[2] - b[2];
    return sqrtf(fmaf(x, x, fmaf(y, y, z * z)));
}


// Interpreter state + resumable execution: run(ops, pc0, pc1, ...) executes ops [pc0, pc1) and keeps the
// kinematic state, so a caller that receives the quad values in pieces can interleave (tensor-core kernel).
// RESID: residual-vector form (reference motion_primitive_constraints.py:124-146 + the constraint classes'
// get_residual_vector_spline): the `term` handed to emit() is a RESIDUAL index taken from the residual program
// (mgb_constraints::ops_resid); the two-hands term emits its 3 distances, the pose term its per-point distances (no mean,
// no velocity term), a local trajectory one error per frame.
template <int CPT, bool RESID = false>
struct MgbInterp {
    MgbXf G[CPT];
    float al_c[CPT], al_s[CPT], al_dx[CPT], al_dz[CPT], al_qw[CPT], al_qy[CPT];
    float scratch[CPT][MGB_SCRATCH_FLOATS];
    float stack[CPT][MGB_STACK_DEPTH * 8];
    int sp;

    __device__ __forceinline__ void init() {
        sp = 0;
#pragma unroll
        for (int c = 0; c < CPT; ++c) {
            mgb_xf_identity(G[c]);
            al_c[c] = 1.f; al_s[c] = 0.f; al_dx[c] = 0.f; al_dz[c] = 0.f; al_qw[c] = 1.f; al_qy[c] = 0.f;
        }
    }

    template <typename QuadFn, typename EmitFn>
    __device__ __forceinline__ void run(const MgbOp* __restrict__ ops, int pc0, int pc1, const float* __restrict__ cst, QuadFn quad, EmitFn emit) {
        if (pc0 >= pc1) return;
        MgbOp next = ops[pc0];
        for (int pc = pc0; pc < pc1; ++pc) {
            const MgbOp op = next;
            if (pc + 1 < pc1) next = ops[pc + 1];   // the decode of the next op overlaps this op's arithmetic
            const float* cs = cst + op.c;
            switch (op.code) {
                case OP_ROOT: {
                    float4 tr[CPT], q[CPT];
                    quad(op.a, tr);
                    quad(op.b, q);
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        // optional alignment to the previous motion (identity when local)
                        const float tx = al_c[c] * tr[c].x + al_s[c] * tr[c].z + al_dx[c];
                        const float tz = -al_s[c] * tr[c].x + al_c[c] * tr[c].z + al_dz[c];
                        const float qw = al_qw[c] * q[c].x - al_qy[c] * q[c].z;
                        const float qx = al_qw[c] * q[c].y + al_qy[c] * q[c].w;
                        const float qy = al_qw[c] * q[c].z + al_qy[c] * q[c].x;
                        const float qz = al_qw[c] * q[c].w - al_qy[c] * q[c].y;
                        mgb_xf_set_rot(G[c], qw, qx, qy, qz);
                        G[c].tx = tx + cs[0];
                        G[c].ty = tr[c].y + cs[1];
                        G[c].tz = tz + cs[2];
                    }
                    break;
                }
                case OP_JOINT: {
                    float4 q[CPT];
                    quad(op.a, q);
#pragma unroll
                    for (int c = 0; c < CPT; ++c) mgb_xf_child(G[c], q[c].x, q[c].y, q[c].z, q[c].w, cs[0], cs[1], cs[2]);
                    break;
                }
                case OP_CONST: {
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        if (op.a) mgb_xf_translate(G[c], cs[4], cs[5], cs[6]);
                        else mgb_xf_child(G[c], cs[0], cs[1], cs[2], cs[3], cs[4], cs[5], cs[6]);
                    }
                    break;
                }
                case OP_PUSH: {
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        float* st = stack[c] + sp * 8;
                        st[0] = G[c].w; st[1] = G[c].x; st[2] = G[c].y; st[3] = G[c].z;
                        st[4] = G[c].s2; st[5] = G[c].tx; st[6] = G[c].ty; st[7] = G[c].tz;
                    }
                    ++sp;
                    break;
                }
                case OP_POP: {
                    --sp;
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        const float* st = stack[c] + sp * 8;
                        G[c].w = st[0]; G[c].x = st[1]; G[c].y = st[2]; G[c].z = st[3];
                        G[c].s2 = st[4]; G[c].tx = st[5]; G[c].ty = st[6]; G[c].tz = st[7];
                    }
                    break;
                }
                case OP_ALIGN: {
                    // heading of the aligning node at control point 0, raw root translation of
                    // control point 0 (quad op.a), previous heading / root
                    float4 tr[CPT];
                    quad(op.a, tr);
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        float px, py, pz;
                        mgb_xf_rotate(G[c], cs[0], cs[1], cs[2], px, py, pz);
                        float ang = cs[3] - atan2f(px, pz);
                        const float PI = 3.14159265358979323846f;
                        if (ang > PI) ang -= 2.f * PI;
                        if (ang <= -PI) ang += 2.f * PI;
                        sincosf(ang, &al_s[c], &al_c[c]);
                        sincosf(0.5f * ang, &al_qy[c], &al_qw[c]);
                        al_dx[c] = cs[4] - (al_c[c] * tr[c].x + al_s[c] * tr[c].z);
                        al_dz[c] = cs[5] - (-al_s[c] * tr[c].x + al_c[c] * tr[c].z);
                    }
                    break;
                }
                case OP_POS: {
#pragma unroll
                    for (int c = 0; c < CPT; ++c) emit(c, op.a, cs[0], masked_dist((unsigned)op.b, cs + 1, G[c].tx, G[c].ty, G[c].tz));
                    break;
                }
                case OP_ORIENT: {
                    // angle_between_vectors(v1, v2) = acos(v1.v2 / |v1||v2|), evaluated as
                    // atan2(|v1 x v2|, v1.v2): same value, but well conditioned near 0 and pi in fp32
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        float vx, vy, vz;
                        mgb_xf_rotate(G[c], 0.f, 0.f, 1.f, vx, vy, vz);
                        const float cx = vy * cs[3] - vz * cs[2], cy = vz * cs[1] - vx * cs[3], cz = vx * cs[2] - vy * cs[1];
                        const float d = vx * cs[1] + vy * cs[2] + vz * cs[3];
                        emit(c, op.a, cs[0], atan2f(sqrtf(cx * cx + cy * cy + cz * cz), d));
                    }
                    break;
                }
                case OP_REL: {
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        float pp[3];
                        mgb_xf_rotate(G[c], cs[1], cs[2], cs[3], pp[0], pp[1], pp[2]);
                        pp[0] = fmaf(G[c].tx, cs[4], pp[0]);
                        pp[1] = fmaf(G[c].ty, cs[4], pp[1]);
                        pp[2] = fmaf(G[c].tz, cs[4], pp[2]);
                        emit(c, op.a, cs[0], dist3(cs + 5, pp));
                    }
                    break;
                }
                case OP_DIR2D: {
                    // degrees(acos(clamp(t.d / (|t||d|)))) evaluated as atan2(|t x d|, t.d) (same value,
                    // well conditioned for small angles in fp32); d = (x, z) of R_global * ref_dir
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        float px, py, pz;
                        mgb_xf_rotate(G[c], cs[1], cs[2], cs[3], px, py, pz);
                        const float cr = cs[4] * pz - cs[5] * px;
                        const float dt = cs[4] * px + cs[5] * pz;
                        emit(c, op.a, cs[0], atan2f(fabsf(cr), dt) * 57.29577951308232f);
                    }
                    break;
                }
                case OP_LOOKAT: {
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        float tx = cs[1] - G[c].tx, ty = cs[2] - G[c].ty, tz = cs[3] - G[c].tz;
                        const float ti = rsqrtf(tx * tx + ty * ty + tz * tz);
                        tx *= ti; ty *= ti; tz *= ti;
                        float hx, hy, hz;
                        mgb_xf_rotate(G[c], 0.f, 0.f, 1.f, hx, hy, hz);
                        const float hi = rsqrtf(hx * hx + hy * hy + hz * hz);
                        hx *= hi; hy *= hi; hz *= hi;
                        const float cx = hy * tz - hz * ty, cy = hz * tx - hx * tz, cz = hx * ty - hy * tx;
                        emit(c, op.a, cs[0], fabsf(atan2f(sqrtf(cx * cx + cy * cy + cz * cz), hx * tx + hy * ty + hz * tz)));
                    }
                    break;
                }
                case OP_STORE: {
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        scratch[c][op.a] = G[c].tx;
                        scratch[c][op.a + 1] = G[c].ty;
                        scratch[c][op.a + 2] = G[c].tz;
                    }
                    break;
                }
                case OP_TWOHANDS: {
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        const float* lp = scratch[c] + op.b;
                        const float* rp = scratch[c] + op.b + 3;
                        const float mid[3] = {lp[0] + 0.5f * (rp[0] - lp[0]), lp[1] + 0.5f * (rp[1] - lp[1]), lp[2] + 0.5f * (rp[2] - lp[2])};
                        if (RESID) {   // two_hand_constraint.py:68-80
                            emit(c, op.a, cs[0], dist3(cs + 1, mid));
                            emit(c, op.a + 1, cs[0], dist3(cs + 4, lp));
                            emit(c, op.a + 2, cs[0], dist3(cs + 7, rp));
                        } else {
                            emit(c, op.a, cs[0], dist3(cs + 1, mid) + dist3(cs + 4, lp) + dist3(cs + 7, rp));
                        }
                    }
                    break;
                }
                case OP_FEET: {
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        const float* lp = scratch[c] + op.b;
                        const float* rp = scratch[c] + op.b + 3;
                        emit(c, op.a, cs[0], (dist3(cs + 1, lp) + dist3(cs + 4, rp)) * cs[7]);
                    }
                    break;
                }
                case OP_POSE: {
                    // weighted 2-D rigid alignment of the candidate's point cloud (b) onto the
                    // target cloud (a), then mean point distance + root velocity error
                    const int np = (int)cs[1];
                    const float W = cs[6], sax = cs[7], saz = cs[8];
                    const float* tg = cs + 9;  // (ax, ay, az, w) * np
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        const float* pts = scratch[c] + op.b;
                        float num_l = 0.f, den_l = 0.f, sbx = 0.f, sbz = 0.f;
                        for (int j = 0; j < np; ++j) {
                            const float ax = tg[4 * j], az = tg[4 * j + 2], w = tg[4 * j + 3];
                            const float bx = pts[3 * j], bz = pts[3 * j + 2];
                            num_l = fmaf(w, ax * bz - bx * az, num_l);
                            den_l = fmaf(w, ax * bx + az * bz, den_l);
                            sbx = fmaf(w, bx, sbx);
                            sbz = fmaf(w, bz, sbz);
                        }
                        const float invW = 1.f / W;
                        const float num = num_l - (sax * sbz - sbx * saz) * invW;
                        const float den = den_l - (sax * sbx + saz * sbz) * invW;
                        const float theta = atan2f(num, den);
                        float st, ct;
                        sincosf(theta, &st, &ct);
                        const float ox = (sax - sbx * ct - sbz * st) * invW;
                        const float oz = (saz + sbx * st - sbz * ct) * invW;
                        float dsum = 0.f;
                        for (int j = 0; j < np; ++j) {
                            const float bx = pts[3 * j], by = pts[3 * j + 1], bz = pts[3 * j + 2];
                            const float tx = ct * bx + st * bz + ox;
                            const float tz = -st * bx + ct * bz + oz;
                            const float ex = tg[4 * j] - tx, ey = tg[4 * j + 1] - by, ez = tg[4 * j + 2] - tz;
                            const float dj = sqrtf(fmaf(ex, ex, fmaf(ey, ey, ez * ez)));
                            dsum += dj;
                            if (RESID) emit(c, op.a + j, cs[0], dj);   // pose_constraint.py:107-121
                        }
                        if (RESID) continue;
                        float e = dsum / (float)np;
                        if (cs[2] != 0.f) {
                            const float* r2 = pts + 3 * np;
                            const float vx = cs[3] - (r2[0] - pts[0]), vy = cs[4] - (r2[1] - pts[1]), vz = cs[5] - (r2[2] - pts[2]);
                            e += sqrtf(fmaf(vx, vx, fmaf(vy, vy, vz * vz)));
                        }
                        emit(c, op.a, cs[0], e);
                    }
                    break;
                }
                case OP_CA_BEGIN: {
#pragma unroll
                    for (int c = 0; c < CPT; ++c) scratch[c][op.a] = INFINITY;
                    break;
                }
                case OP_CA_ACC: {
#pragma unroll
                    for (int c = 0; c < CPT; ++c)
                        scratch[c][op.a] = fminf(scratch[c][op.a], masked_dist((unsigned)op.b, cs, G[c].tx, G[c].ty, G[c].tz));
                    break;
                }
                case OP_CA_END: {
#pragma unroll
                    for (int c = 0; c < CPT; ++c) emit(c, op.a, cs[0], scratch[c][op.b]);
                    break;
                }
                case OP_JROT: {
                    // |R_target - R(q / |q|)|_F of one joint's own quaternion channels (no FK): JointRotationConstraint
                    // (joint_rotation_constraint.py:62-75) and, summed with inner weights, PoseConstraintQuatFrame
                    float4 q[CPT];
                    quad(op.b, q);
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        float qw = q[c].x, qx = q[c].y, qy = q[c].z, qz = q[c].w;
                        if (cs[11] != 0.f) {   // root joint: the aligned frame carries the aligning rotation about +y
                            const float w2 = al_qw[c] * qw - al_qy[c] * qy, x2 = al_qw[c] * qx + al_qy[c] * qz;
                            const float y2 = al_qw[c] * qy + al_qy[c] * qw, z2 = al_qw[c] * qz - al_qy[c] * qx;
                            qw = w2; qx = x2; qy = y2; qz = z2;
                        }
                        const float n = fmaf(qw, qw, fmaf(qx, qx, fmaf(qy, qy, qz * qz)));
                        const float s2 = __fdividef(2.f, n);
                        const float xx = qx * qx * s2, yy = qy * qy * s2, zz = qz * qz * s2;
                        const float xy = qx * qy * s2, xz = qx * qz * s2, yz = qy * qz * s2;
                        const float wx = qw * qx * s2, wy = qw * qy * s2, wz = qw * qz * s2;
                        const float R[9] = {1.f - yy - zz, xy - wz, xz + wy, xy + wz, 1.f - xx - zz, yz - wx, xz - wy, yz + wx, 1.f - xx - yy};
                        float d = 0.f;
#pragma unroll
                        for (int i = 0; i < 9; ++i) { const float e = cs[2 + i] - R[i]; d = fmaf(e, e, d); }
                        emit(c, op.a, cs[0], cs[1] * sqrtf(d));
                    }
                    break;
                }
                case OP_LT_BEGIN: {
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        scratch[c][op.a] = 0.f;
                        scratch[c][op.a + 1] = cs[0];
                        scratch[c][op.a + 5] = 0.f;
                    }
                    break;
                }
                case OP_LT_ACC: {
                    // LocalTrajectoryConstraint.get_spline_distance (local_trajectory_constraint.py:66-83): the arc length walked by
                    // the joint so far picks the target on the trajectory, tabulated by arc length (linear between samples)
                    const float step_inv = cs[0], full = cs[2];
                    const int nt = (int)cs[1];
                    const float* tb = cs + 4;   // cs[3] = weight (residual form)
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        float* st = scratch[c] + op.a;
                        const float px = G[c].tx, py = G[c].ty, pz = G[c].tz;
                        if (st[5] != 0.f) {
                            const float dx = st[2] - px, dy = st[3] - py, dz = st[4] - pz;
                            st[1] += sqrtf(fmaf(dx, dx, fmaf(dy, dy, dz * dz)));
                        }
                        float tx, tz;
                        if (st[1] <= full) {
                            const float u = fmaxf(st[1], 0.f) * step_inv;
                            int i0 = min((int)u, nt - 1);
                            const float f = u - (float)i0;
                            tx = fmaf(f, tb[2 * i0 + 2] - tb[2 * i0], tb[2 * i0]);
                            tz = fmaf(f, tb[2 * i0 + 3] - tb[2 * i0 + 1], tb[2 * i0 + 1]);
                        } else {   // beyond the end: the last control point
                            tx = tb[2 * nt + 2];
                            tz = tb[2 * nt + 3];
                        }
                        const float ex = tx - px, ez = tz - pz;
                        st[0] += fmaf(ex, ex, ez * ez);
                        if (RESID) emit(c, op.b, tb[-1], fmaf(ex, ex, ez * ez));   // one residual per frame; tb[-1] = weight
                        st[2] = px; st[3] = py; st[4] = pz; st[5] = 1.f;
                    }
                    break;
                }
                case OP_LT_END: {
                    if (RESID) break;
#pragma unroll
                    for (int c = 0; c < CPT; ++c) emit(c, op.a, cs[0], scratch[c][op.b]);
                    break;
                }
                case OP_SCR_INIT: {
                    for (int i = 0; i < op.b; ++i) {
                        const float v = cs[i];
#pragma unroll
                        for (int c = 0; c < CPT; ++c) scratch[c][op.a + i] = v;
                    }
                    break;
                }
                case OP_TR_ACC: {
                    // TrajectoryConstraint.get_residual_vector (trajectory_constraint.py:95-116): the closest point of the curve at
                    // a parameter >= the previous frame's.  The reference runs L-BFGS-B on |S(u) - p| from the lower bound
                    // (parameterized_spline.py:303-322); here the curve is a polyline over n_table equal parameter steps and the
                    // walk goes up from the bound until the distance rises again (the first local minimum).
                    const int nt = (int)cs[0];
                    const float* tb = cs + 4;
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        float* st = scratch[c] + op.a;
                        const float px = G[c].tx, py = G[c].ty, pz = G[c].tz;
                        const float x0 = fminf(fmaxf(st[1], 0.f), 1.f) * (float)nt;
                        const int i0 = min((int)x0, nt - 1);
                        float lo = x0 - (float)i0;
                        float best = 0.f, best_x = x0;
                        // walk up while the distance still falls at the END of the segment, i.e. while the foot of the perpendicular
                        // lies beyond it (decided on the projection, not by comparing squared distances: far from the curve
                        // their differences drown in float32 rounding)
                        for (int i = i0; i < nt; ++i) {
                            const float ax = tb[3 * i], ay = tb[3 * i + 1], az = tb[3 * i + 2];
                            const float dx = tb[3 * i + 3] - ax, dy = tb[3 * i + 4] - ay, dz = tb[3 * i + 5] - az;
                            const float ex = px - ax, ey = py - ay, ez = pz - az;
                            const float dd = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
                            const float pr = fmaf(ex, dx, fmaf(ey, dy, ez * dz));
                            if (pr >= dd && i + 1 < nt) {   // includes dd == 0 with pr == 0: a repeated point, step over it
                                lo = 0.f;
                                continue;
                            }
                            const float sgm = fminf(fmaxf(dd > 0.f ? __fdividef(pr, dd) : lo, lo), 1.f);
                            const float rx = fmaf(-sgm, dx, ex), ry = fmaf(-sgm, dy, ey), rz = fmaf(-sgm, dz, ez);
                            best = fmaf(rx, rx, fmaf(ry, ry, rz * rz));
                            best_x = (float)i + sgm;
                            break;
                        }
                        const float e = sqrtf(best);
                        st[0] += e;
                        st[1] = best_x / (float)nt;
                        if (RESID) emit(c, op.b, cs[1], e);
                    }
                    break;
                }
                case OP_TS_ACC: {
                    // TrajectorySetConstraint.get_residual_vector (trajectory_set_constraint.py:83-105): every joint walks its own
                    // trajectory by the arc length of its own path; a frame is scored with the arc lengths reached BEFORE it
                    const float step_inv = cs[0], full = cs[2], r0 = cs[3], r1 = cs[4];
                    const int nt = (int)cs[1];
                    const float* tb = cs + 8;
                    const int jj = (op.b >> 1) & 127;
#pragma unroll
                    for (int c = 0; c < CPT; ++c) {
                        float* st = scratch[c] + op.a;
                        float* js = st + 3 + 5 * jj;
                        const float px = G[c].tx, py = G[c].ty, pz = G[c].tz;
                        const float arc = js[0];
                        if (arc >= r0 && arc <= r1) st[2] = 1.f;
                        float tx, ty, tz;
                        if (arc <= full) {
                            const float u = fmaxf(arc, 0.f) * step_inv;
                            const int i0 = min((int)u, nt - 1);
                            const float f = u - (float)i0;
                            tx = fmaf(f, tb[3 * i0 + 3] - tb[3 * i0], tb[3 * i0]);
                            ty = fmaf(f, tb[3 * i0 + 4] - tb[3 * i0 + 1], tb[3 * i0 + 1]);
                            tz = fmaf(f, tb[3 * i0 + 5] - tb[3 * i0 + 2], tb[3 * i0 + 2]);
                        } else {
                            tx = tb[3 * nt + 3]; ty = tb[3 * nt + 4]; tz = tb[3 * nt + 5];
                        }
                        st[1] += (px - tx) + (py - ty) + (pz - tz);
                        if (js[4] != 0.f) {
                            const float dx = js[1] - px, dy = js[2] - py, dz = js[3] - pz;
                            js[0] = arc + sqrtf(fmaf(dx, dx, fmaf(dy, dy, dz * dz)));
                        }
                        js[1] = px; js[2] = py; js[3] = pz; js[4] = 1.f;
                        if (op.b & 1) {   // last joint of the frame
                            const float e = st[2] != 0.f ? fabsf(st[1]) * cs[5] : 0.f;
                            st[0] += e;
                            st[1] = 0.f;
                            st[2] = 0.f;
                            if (RESID) emit(c, op.b >> 8, cs[6], e);
                        }
                    }
                    break;
                }
                case OP_MEAN_END: {
                    if (RESID) break;
#pragma unroll
                    for (int c = 0; c < CPT; ++c) emit(c, op.a, cs[0], scratch[c][op.b] * cs[1]);
                    break;
                }
                default:
                    break;
            }
        }

    }
};

template <int CPT, typename QuadFn, typename EmitFn>
__device__ __forceinline__ void mgb_run_program(const MgbOp* __restrict__ ops, int n_ops, const float* __restrict__ cst, QuadFn quad, EmitFn emit) {
    MgbInterp<CPT> it;
    it.init();
    it.run(ops, 0, n_ops, cst, quad, emit);
}
