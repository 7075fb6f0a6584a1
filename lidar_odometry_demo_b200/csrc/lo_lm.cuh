// K6: the trust-region Levenberg-Marquardt iteration that replaces ceres::Solve for the program
// CloudMatcher::align builds (src/cloud_matcher.cpp:109-158), on the 6x6 normal equations.
//
// What is reproduced (Ceres 2.x TrustRegionMinimizer + LevenbergMarquardtStrategy, options of
// cloud_matcher.cpp:110-112 plus defaults): Jacobi column scaling fixed at iteration 0,
// diagonal clamp [1e-6,1e32], D = sqrt(diag/radius), step = -(Hs + D^2)^-1 gs (Ceres solves the same
// least-squares problem by dense QR of [J;D]), model cost change, step validity, candidate = Plus(x, step*scale)
// with QuaternionManifold::Plus, parameter / function tolerance exits that do not apply the step, acceptance
// on rho > 1e-3, radius update, at most 4 iterations, and the step_norm of the last RECORDED iteration that
// the outer loop tests (cloud_matcher.cpp:169). Everything is double. The state is a plain struct so that
// every thread block can run the (tiny, deterministic) scalar logic redundantly on identical inputs.
#pragma once

#include "lo_match.cuh"

namespace lo {

enum LmTermination { LM_NO_CONVERGENCE = 0, LM_CONV_FUNCTION = 1, LM_CONV_PARAMETER = 2, LM_CONV_GRADIENT = 3, LM_FAILURE = 4, LM_CONV_RADIUS = 5 };

struct LmState {
    double x[7];           // qw,qx,qy,qz,tx,ty,tz
    double x_norm, x_cost;
    double H[21];          // upper triangle of JtJ at x (unscaled, prior included)
    double g[6];           // Jtr at x
    double scale[6];       // Jacobi scaling, fixed at iteration 0
    double diagonal[6];
    double radius, decrease_factor;
    double cand[7];
    double model_cost_change;
    double last_step_norm; // step_norm of the last recorded iteration
    double grad_max;
    int reuse_diagonal, iteration, num_invalid, any_success, termination, recorded;
};

__host__ __device__ inline int tri(int a, int b) { return a * 6 - (a * (a - 1)) / 2 + (b - a); }  // a <= b
__host__ __device__ inline double sym(const double* H, int a, int b) { return a <= b ? H[tri(a, b)] : H[tri(b, a)]; }

// NormalPrior(diag(10), b) with TrivialLoss (cloud_matcher.cpp:153-154): three rows 10*(t-b), Jacobian 10*I
__host__ __device__ inline void add_prior(double acc[kAcc], const double t[3], const double b[3]) {
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const double r = 10.0 * (t[k] - b[k]);
        acc[27] += 0.5 * (r * r);
        acc[21 + 3 + k] += 10.0 * r;
        acc[tri(3 + k, 3 + k)] += 100.0;
    }
}

__host__ __device__ inline void quaternion_plus(const double x[4], const double d[3], double out[4]) {
    const double nd = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    if (nd == 0.0) {
#pragma unroll
        for (int i = 0; i < 4; i++) out[i] = x[i];
        return;
    }
    double sn, q0;
    sincos(nd, &sn, &q0);
    const double s = sn / nd;
    const double q1 = s * d[0], q2 = s * d[1], q3 = s * d[2];
    out[0] = q0 * x[0] - q1 * x[1] - q2 * x[2] - q3 * x[3];
    out[1] = q0 * x[1] + q1 * x[0] + q2 * x[3] - q3 * x[2];
    out[2] = q0 * x[2] - q1 * x[3] + q2 * x[0] + q3 * x[1];
    out[3] = q0 * x[3] + q1 * x[2] - q2 * x[1] + q3 * x[0];
}

__host__ __device__ inline double norm7(const double v[7]) {
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 7; i++) s += v[i] * v[i];
    return sqrt(s);
}

// FinalizeIterationAndCheckIfMinimizerCanContinue: record the iteration, then the stop tests
__host__ __device__ inline bool lm_record(LmState& s, double step_norm) {
    s.recorded++;
    s.last_step_norm = step_norm;
    if (s.iteration >= 4) { s.termination = LM_NO_CONVERGENCE; return false; }
    if (s.grad_max <= 1e-10) { s.termination = LM_CONV_GRADIENT; return false; }
    if (s.radius <= 1e-32) { s.termination = LM_CONV_RADIUS; return false; }
    return true;
}

__host__ __device__ inline void lm_take_point(LmState& s, const double acc[kAcc]) {
#pragma unroll
    for (int k = 0; k < 21; k++) s.H[k] = acc[k];
#pragma unroll
    for (int k = 0; k < 6; k++) s.g[k] = acc[21 + k];
    s.x_cost = acc[27];
    // gradient_max_norm as Ceres computes it: the infinity norm of x - Plus(x, -g) over the 7 ambient coordinates
    // (TrustRegionMinimizer::EvaluateGradientAndJacobian); s.x is the point the sums were taken at
    const double ng[3] = {-s.g[0], -s.g[1], -s.g[2]};
    double qp[4];
    quaternion_plus(s.x, ng, qp);
    double gm = 0.0;
#pragma unroll
    for (int k = 0; k < 4; k++) gm = fmax(gm, fabs(s.x[k] - qp[k]));
#pragma unroll
    for (int k = 0; k < 3; k++) gm = fmax(gm, fabs(s.x[4 + k] - (s.x[4 + k] - s.g[3 + k])));
    s.grad_max = gm;
}

// IterationZero. acc = normal equations at x0 (prior included). Returns true if the solve continues.
__host__ __device__ inline bool lm_begin(LmState& s, const double x0[7], const double acc[kAcc]) {
#pragma unroll
    for (int i = 0; i < 7; i++) s.x[i] = x0[i];
    s.x_norm = norm7(s.x);
    lm_take_point(s, acc);
#pragma unroll
    for (int k = 0; k < 6; k++) s.scale[k] = 1.0 / (1.0 + sqrt(s.H[tri(k, k)]));
    s.radius = 1e4;
    s.decrease_factor = 2.0;
    s.reuse_diagonal = 0;
    s.iteration = 0;
    s.num_invalid = 0;
    s.any_success = 0;
    s.termination = LM_NO_CONVERGENCE;
    s.recorded = 0;
    return lm_record(s, 0.0);
}

// Solve (Hs + diag(d2)) y = gs by the square-root-free Cholesky factorisation A = L D L' (unit lower L); false if A is
// not positive definite or the solution is not finite. One reciprocal per column is the only long-latency operation
// on the dependent chain (the whole step logic runs on one thread between two evaluations).
__host__ __device__ inline bool solve6(const double Hs[21], const double d2[6], const double gs[6], double y[6]) {
    double L[6][6], LD[6][6], inv[6];   // LD[a][k] = L[a][k] * D[k]
#pragma unroll
    for (int j = 0; j < 6; j++) {
        double d = Hs[tri(j, j)] + d2[j];
#pragma unroll
        for (int k = 0; k < j; k++) d = fma(-LD[j][k], L[j][k], d);
        if (!(d > 0.0)) return false;
        inv[j] = 1.0 / d;
#pragma unroll
        for (int a = j + 1; a < 6; a++) {
            double v = Hs[tri(j, a)];
#pragma unroll
            for (int k = 0; k < j; k++) v = fma(-LD[a][k], L[j][k], v);
            LD[a][j] = v;
            L[a][j] = v * inv[j];
        }
    }
    double z[6];
#pragma unroll
    for (int a = 0; a < 6; a++) {
        double v = gs[a];
#pragma unroll
        for (int k = 0; k < a; k++) v = fma(-L[a][k], z[k], v);
        z[a] = v;
    }
#pragma unroll
    for (int a = 5; a >= 0; a--) {
        double v = z[a] * inv[a];
#pragma unroll
        for (int k = a + 1; k < 6; k++) v = fma(-L[k][a], y[k], v);
        y[a] = v;
    }
#pragma unroll
    for (int a = 0; a < 6; a++)
        if (!isfinite(y[a])) return false;
    return true;
}

// ComputeTrustRegionStep (+ HandleInvalidStep). Returns true when s.cand holds a candidate that must be
// evaluated; false when the solve is over (s.termination set).
__host__ __device__ inline bool lm_propose(LmState& s) {
    while (true) {
        double Hs[21], gs[6];
#pragma unroll
        for (int a = 0; a < 6; a++) {
            gs[a] = s.g[a] * s.scale[a];
#pragma unroll
            for (int b = a; b < 6; b++) Hs[tri(a, b)] = s.H[tri(a, b)] * s.scale[a] * s.scale[b];
        }
        if (!s.reuse_diagonal)
#pragma unroll
            for (int a = 0; a < 6; a++) s.diagonal[a] = fmin(fmax(Hs[tri(a, a)], 1e-6), 1e32);
        double d2[6], y[6], step[6];
        const double inv_radius = 1.0 / s.radius;
#pragma unroll
        for (int a = 0; a < 6; a++) d2[a] = s.diagonal[a] * inv_radius;
        const bool ok = solve6(Hs, d2, gs, y);
        s.reuse_diagonal = 1;
        bool valid = false;
        if (ok) {
#pragma unroll
            for (int a = 0; a < 6; a++) step[a] = -y[a];
            // model_cost_change = -(J s)'(r + J s / 2) = -(s'gs + s'Hs s / 2)
            double lin = 0.0, quad = 0.0;
#pragma unroll
            for (int a = 0; a < 6; a++) {
                lin += step[a] * gs[a];
                double row = 0.0;
#pragma unroll
                for (int b = 0; b < 6; b++) row += sym(Hs, a, b) * step[b];
                quad += step[a] * row;
            }
            s.model_cost_change = -(lin + 0.5 * quad);
            valid = s.model_cost_change > 0.0;
        }
        s.iteration += 1;
        if (valid) {
            s.num_invalid = 0;
            double delta[6];
#pragma unroll
            for (int a = 0; a < 6; a++) delta[a] = step[a] * s.scale[a];
            quaternion_plus(s.x, delta, s.cand);
#pragma unroll
            for (int k = 0; k < 3; k++) s.cand[4 + k] = s.x[4 + k] + delta[3 + k];
            return true;
        }
        if (++s.num_invalid >= 5) {
            s.termination = LM_FAILURE;
            return false;
        }
        s.radius = s.radius / s.decrease_factor;
        s.decrease_factor *= 2.0;
        if (!lm_record(s, 0.0)) return false;
    }
}

// Everything after ComputeCandidatePointAndEvaluateCost. acc = normal equations at s.cand (prior included).
// Returns true if another lm_propose should follow.
__host__ __device__ inline bool lm_update(LmState& s, const double acc[kAcc]) {
    double cand_cost = acc[27];
    if (!isfinite(cand_cost)) cand_cost = 1.7976931348623157e308;
    double step_norm = 0.0;
    if (s.any_success) {  // Ceres 2.2: parameter tolerance only after one successful step
        double d = 0.0;
#pragma unroll
        for (int i = 0; i < 7; i++) d += (s.x[i] - s.cand[i]) * (s.x[i] - s.cand[i]);
        step_norm = sqrt(d);
        if (step_norm <= 1e-8 * (s.x_norm + 1e-8)) { s.termination = LM_CONV_PARAMETER; return false; }
    }
    const double cost_change = s.x_cost - cand_cost;
    if (fabs(cost_change) <= 1e-5 * s.x_cost) { s.termination = LM_CONV_FUNCTION; return false; }
    const double rho = cand_cost >= 1.7976931348623157e308 ? -1.7976931348623157e308 : cost_change / s.model_cost_change;
    if (rho > 1e-3) {
#pragma unroll
        for (int i = 0; i < 7; i++) s.x[i] = s.cand[i];
        s.x_norm = norm7(s.x);
        lm_take_point(s, acc);
        const double t = 2.0 * rho - 1.0;
        s.radius = s.radius / fmax(1.0 / 3.0, 1.0 - t * t * t);
        s.radius = fmin(1e16, s.radius);
        s.decrease_factor = 2.0;
        s.reuse_diagonal = 0;
        s.any_success = 1;
    } else {
        s.radius = s.radius / s.decrease_factor;
        s.decrease_factor *= 2.0;
        s.reuse_diagonal = 1;
    }
    return lm_record(s, step_norm);
}

}  // namespace lo
