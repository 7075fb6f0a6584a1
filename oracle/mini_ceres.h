// TEST INFRASTRUCTURE (oracle/). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may build or call this; the product (lidar_odometry_demo_b200/) never does.
//
// mini_ceres: a CPU restatement of what `ceres::Solve` does for the one problem the reference builds in
// CloudMatcher::align (/root/reference/src/cloud_matcher.cpp:109-158):
//   parameter blocks  quat[4] (w,x,y,z) with ceres::QuaternionManifold, translation[3]
//   residual blocks   N x PointToPlaneErrorAnalytic (SizedCostFunction<1,4,3>) with HuberLoss(0.15)
//                     1 x NormalPrior(diag(10,10,10), guess.t) with TrivialLoss
//   options           DENSE_QR, max_num_iterations = 4, function_tolerance = 1e-5, rest default
//
// PARITY UNPINNED: Ceres Solver is a third-party dependency that is NOT under /root/reference and is
// not installed in this image; the reference does not even pin a version (Dockerfile:8-10 clones HEAD,
// comment says 2.2). This file restates the published algorithm of Ceres 2.1/2.2
// (internal/ceres/trust_region_minimizer.cc, levenberg_marquardt_strategy.cc, dense_qr_solver.cc,
// residual_block.cc, corrector.cc, loss_function.cc, manifold.cc/QuaternionPlusImpl, normal_prior.cc,
// trust_region_step_evaluator.cc) from knowledge of that code. It could not be diffed against a Ceres
// build. The reference's only test of this path is CloudMatcher.MatchingTest (test/test.cpp:191-264,
// 0.05 m / 0.01), which the oracle satisfies on the substitute fixture (tests/test_oracle_cpu.py::test_golden_config1).
#ifndef LO_ORACLE_MINI_CERES_H
#define LO_ORACLE_MINI_CERES_H

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <limits>
#include <vector>

namespace mini_ceres {

struct Pair {        // one VoxelGrid::Correspondence (voxel_grid.h:40-46), already widened to double
    double p[3];     // source_point_local
    double o[3];     // plane_origin
    double n[3];     // plane_normal
};

struct Options {     // cloud_matcher.cpp:109-112 + Ceres 2.x Solver::Options defaults
    int max_num_iterations = 4;
    double function_tolerance = 1e-5;
    double gradient_tolerance = 1e-10;
    double parameter_tolerance = 1e-8;
    double initial_trust_region_radius = 1e4;
    double max_trust_region_radius = 1e16;
    double min_trust_region_radius = 1e-32;
    double min_relative_decrease = 1e-3;
    double min_lm_diagonal = 1e-6;
    double max_lm_diagonal = 1e32;
    int max_num_consecutive_invalid_steps = 5;
    bool jacobi_scaling = true;
    double huber_a = 0.15;       // cloud_matcher.cpp:134
    double prior_weight = 10.0;  // 1/0.1, cloud_matcher.cpp:153 (1.0/0.1 rounds to exactly 10.0)
    // Ceres 2.2 only tests the parameter tolerance after at least one successful step
    // (trust_region_minimizer.cc `atleast_one_successful_step`); step_norm is computed inside that test.
    bool param_tol_needs_successful_step = true;
};

enum Termination { NO_CONVERGENCE = 0, CONV_FUNCTION = 1, CONV_PARAMETER = 2, CONV_GRADIENT = 3, FAILURE = 4, CONV_RADIUS = 5 };

struct IterationLog {  // the fields of ceres::IterationSummary that matter here
    int iteration = 0;
    bool step_is_valid = false;
    bool step_is_successful = false;
    double cost = 0, cost_change = 0, gradient_max_norm = 0, gradient_norm = 0;
    double step_norm = 0, relative_decrease = 0, trust_region_radius = 0;
};

struct Summary {
    std::vector<IterationLog> iterations;
    Termination termination = NO_CONVERGENCE;
    double initial_cost = 0, final_cost = 0;
    int num_cost_evals = 0, num_jacobian_evals = 0;
};

// ---------------------------------------------------------------------------------------------
// Residual + Jacobian of one point-to-plane block (cloud_matcher.cpp:48-102), all double.
// q = (w,x,y,z) is used as given (not normalised), exactly like the reference.
inline void quat_rotate(const double q[4], const double p[3], double out[3]) {
    // Eigen Quaternion * Vector3: uv = q.vec x p; uv += uv; p + w*uv + q.vec x uv
    const double w = q[0], x = q[1], y = q[2], z = q[3];
    double uv[3] = {y * p[2] - z * p[1], z * p[0] - x * p[2], x * p[1] - y * p[0]};
    uv[0] += uv[0]; uv[1] += uv[1]; uv[2] += uv[2];
    const double c[3] = {y * uv[2] - z * uv[1], z * uv[0] - x * uv[2], x * uv[1] - y * uv[0]};
    for (int i = 0; i < 3; i++) out[i] = (p[i] + w * uv[i]) + c[i];
}

inline double dot3(const double a[3], const double b[3]) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }

inline double point_to_plane(const Pair& c, const double q[4], const double t[3], double* jq /*4 or null*/,
                             double* jt /*3 or null*/) {
    double rp[3];
    quat_rotate(q, c.p, rp);
    const double d[3] = {(rp[0] + t[0]) - c.o[0], (rp[1] + t[1]) - c.o[1], (rp[2] + t[2]) - c.o[2]};
    const double r = dot3(d, c.n);
    if (jq) {
        const double q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
        // the four dR/dq_k matrices of cloud_matcher.cpp:64-86 (each already times 2)
        const double dw[3][3] = {{q0, -q3, q2}, {q3, q0, -q1}, {-q2, q1, q0}};
        const double dx[3][3] = {{q1, q2, q3}, {q2, -q1, -q0}, {q3, q0, -q1}};
        const double dy[3][3] = {{-q2, q1, q0}, {q1, q2, q3}, {-q0, q3, -q2}};
        const double dz[3][3] = {{-q3, -q0, q1}, {q0, -q3, q2}, {q1, q2, q3}};
        const double(*ms[4])[3] = {dw, dx, dy, dz};
        for (int k = 0; k < 4; k++) {
            double v[3];
            for (int i = 0; i < 3; i++)
                v[i] = ((2.0 * ms[k][i][0]) * c.p[0] + (2.0 * ms[k][i][1]) * c.p[1]) + (2.0 * ms[k][i][2]) * c.p[2];
            jq[k] = dot3(v, c.n);
        }
    }
    if (jt) { jt[0] = c.n[0]; jt[1] = c.n[1]; jt[2] = c.n[2]; }
    return r;
}

// QuaternionManifold (Ceres order w,x,y,z): Plus and PlusJacobian (manifold.cc, QuaternionPlusImpl).
inline void quaternion_plus(const double x[4], const double delta[3], double out[4]) {
    const double norm_delta = std::sqrt(delta[0] * delta[0] + delta[1] * delta[1] + delta[2] * delta[2]);
    if (norm_delta == 0.0) {
        for (int i = 0; i < 4; i++) out[i] = x[i];
        return;
    }
    const double s = std::sin(norm_delta) / norm_delta;
    const double qd[4] = {std::cos(norm_delta), s * delta[0], s * delta[1], s * delta[2]};
    out[0] = qd[0] * x[0] - qd[1] * x[1] - qd[2] * x[2] - qd[3] * x[3];
    out[1] = qd[0] * x[1] + qd[1] * x[0] + qd[2] * x[3] - qd[3] * x[2];
    out[2] = qd[0] * x[2] - qd[1] * x[3] + qd[2] * x[0] + qd[3] * x[1];
    out[3] = qd[0] * x[3] + qd[1] * x[2] - qd[2] * x[1] + qd[3] * x[0];
}

inline void quaternion_plus_jacobian(const double x[4], double j[4][3]) {
    j[0][0] = -x[1]; j[0][1] = -x[2]; j[0][2] = -x[3];
    j[1][0] = x[0];  j[1][1] = x[3];  j[1][2] = -x[2];
    j[2][0] = -x[3]; j[2][1] = x[0];  j[2][2] = x[1];
    j[3][0] = x[2];  j[3][1] = -x[1]; j[3][2] = x[0];
}

// ---------------------------------------------------------------------------------------------
// Problem evaluation = ProgramEvaluator + ResidualBlock::Evaluate + Corrector for this program.
// Row i < n: point residual i; rows n..n+2: the prior. Tangent columns: 0..2 rotation, 3..5 translation.
struct Problem {
    const Pair* pairs = nullptr;
    std::size_t n = 0;
    double prior_b[3] = {0, 0, 0};
    const Options* opt = nullptr;

    std::size_t rows() const { return n + 3; }

    // x = [q(4), t(3)]. residuals: rows(); jacobian: rows() x 6 row-major or null.
    void evaluate(const double x[7], double* cost, double* residuals, double* jacobian) const {
        const double* q = x;
        const double* t = x + 4;
        double pj[4][3];
        if (jacobian) quaternion_plus_jacobian(q, pj);
        const double a = opt->huber_a, b = a * a;
        double total = 0.0;
        for (std::size_t i = 0; i < n; i++) {
            double jq[4], jt[3];
            double r = point_to_plane(pairs[i], q, t, jacobian ? jq : nullptr, jacobian ? jt : nullptr);
            const double s = r * r;
            double rho0, rho1;
            if (s > b) {  // HuberLoss::Evaluate, outlier region
                const double rr = std::sqrt(s);
                rho0 = 2.0 * a * rr - b;
                rho1 = std::max(std::numeric_limits<double>::min(), a / rr);
            } else {
                rho0 = s;
                rho1 = 1.0;
            }
            total += 0.5 * rho0;
            // Corrector: rho'' <= 0 for Huber, so residual and Jacobian are scaled by sqrt(rho')
            const double sc = std::sqrt(rho1);
            if (jacobian) {
                double* row = jacobian + 6 * i;
                for (int c = 0; c < 3; c++) {
                    double v = 0.0;
                    for (int k = 0; k < 4; k++) v += jq[k] * pj[k][c];
                    row[c] = sc * v;
                }
                for (int c = 0; c < 3; c++) row[3 + c] = sc * jt[c];
            }
            if (residuals) residuals[i] = sc * r;
        }
        // NormalPrior: r = A (t - b), J = A; TrivialLoss leaves both unchanged.
        for (int k = 0; k < 3; k++) {
            const double r = opt->prior_weight * (t[k] - prior_b[k]);
            total += 0.5 * (r * r);
            if (residuals) residuals[n + k] = r;
            if (jacobian) {
                double* row = jacobian + 6 * (n + k);
                for (int c = 0; c < 6; c++) row[c] = 0.0;
                row[3 + k] = opt->prior_weight;
            }
        }
        *cost = total;
    }

    static void plus(const double x[7], const double delta[6], double out[7]) {
        quaternion_plus(x, delta, out);
        for (int k = 0; k < 3; k++) out[4 + k] = x[4 + k] + delta[3 + k];
    }
};

// Least squares min ||A y - b||^2 by Householder QR (what DENSE_QR does on [J; D], [r; 0]).
// A is m x 6 row-major and is overwritten.
inline bool householder_solve6(std::vector<double>& A, std::vector<double>& b, std::size_t m, double y[6]) {
    const int nc = 6;
    for (int k = 0; k < nc; k++) {
        double norm2 = 0.0;
        for (std::size_t i = k; i < m; i++) norm2 += A[i * nc + k] * A[i * nc + k];
        const double norm = std::sqrt(norm2);
        if (norm == 0.0) return false;
        const double alpha = (A[k * nc + k] > 0.0) ? -norm : norm;
        // v = x - alpha e1 (stored in column k from row k), beta = 2 / v'v
        const double v0 = A[k * nc + k] - alpha;
        const double vtv = norm2 - A[k * nc + k] * A[k * nc + k] + v0 * v0;
        if (vtv == 0.0) { A[k * nc + k] = alpha; continue; }
        A[k * nc + k] = v0;
        for (int c = k + 1; c < nc; c++) {
            double s = 0.0;
            for (std::size_t i = k; i < m; i++) s += A[i * nc + k] * A[i * nc + c];
            s = 2.0 * s / vtv;
            for (std::size_t i = k; i < m; i++) A[i * nc + c] -= s * A[i * nc + k];
        }
        {
            double s = 0.0;
            for (std::size_t i = k; i < m; i++) s += A[i * nc + k] * b[i];
            s = 2.0 * s / vtv;
            for (std::size_t i = k; i < m; i++) b[i] -= s * A[i * nc + k];
        }
        A[k * nc + k] = alpha;  // R(k,k); entries below are the (now unused) reflector
    }
    for (int k = nc - 1; k >= 0; k--) {
        double s = b[k];
        for (int c = k + 1; c < nc; c++) s -= A[k * nc + c] * y[c];
        if (A[k * nc + k] == 0.0) return false;
        y[k] = s / A[k * nc + k];
    }
    for (int k = 0; k < nc; k++)
        if (!std::isfinite(y[k])) return false;
    return true;
}

// TrustRegionMinimizer::Minimize for this program. x = [q(4), t(3)] is updated in place to the
// last accepted point (Ceres copies back the minimum-cost iterate; with monotonic steps that is it).
inline void solve(const Problem& problem, double x_inout[7], Summary* summary) {
    const Options& opt = *problem.opt;
    const std::size_t m = problem.rows();
    const int nc = 6;
    summary->iterations.clear();
    summary->termination = NO_CONVERGENCE;
    summary->num_cost_evals = summary->num_jacobian_evals = 0;

    double x[7];
    for (int i = 0; i < 7; i++) x[i] = x_inout[i];
    double x_norm = 0.0;
    for (int i = 0; i < 7; i++) x_norm += x[i] * x[i];
    x_norm = std::sqrt(x_norm);

    std::vector<double> residuals(m), jac(m * nc), model_res(m);
    double x_cost = 0.0;
    double gradient[6];
    double scale[6] = {1, 1, 1, 1, 1, 1};
    double gradient_max_norm = 0.0, gradient_norm = 0.0;

    // strategy state (LevenbergMarquardtStrategy)
    double radius = opt.initial_trust_region_radius;
    double decrease_factor = 2.0;
    bool reuse_diagonal = false;
    double diagonal[6];

    auto evaluate_gradient_and_jacobian = [&](int iteration) {
        problem.evaluate(x, &x_cost, residuals.data(), jac.data());
        summary->num_jacobian_evals++;
        // gradient = J' r (unscaled Jacobian), tangent space
        for (int c = 0; c < nc; c++) gradient[c] = 0.0;
        for (std::size_t i = 0; i < m; i++)
            for (int c = 0; c < nc; c++) gradient[c] += jac[i * nc + c] * residuals[i];
        // TrustRegionMinimizer::EvaluateGradientAndJacobian: the norms are those of the projected gradient step in the
        // AMBIENT space, x - Plus(x, -gradient) (7 coordinates), not of the tangent-space gradient itself
        gradient_max_norm = 0.0;
        gradient_norm = 0.0;
        {
            double negative_gradient[6], projected[7];
            for (int c = 0; c < nc; c++) negative_gradient[c] = -gradient[c];
            Problem::plus(x, negative_gradient, projected);
            for (int k = 0; k < 7; k++) {
                const double d = x[k] - projected[k];
                gradient_max_norm = std::max(gradient_max_norm, std::fabs(d));
                gradient_norm += d * d;
            }
        }
        gradient_norm = std::sqrt(gradient_norm);
        if (opt.jacobi_scaling) {
            if (iteration == 0) {
                for (int c = 0; c < nc; c++) {
                    double s2 = 0.0;
                    for (std::size_t i = 0; i < m; i++) s2 += jac[i * nc + c] * jac[i * nc + c];
                    scale[c] = 1.0 / (1.0 + std::sqrt(s2));
                }
            }
            for (std::size_t i = 0; i < m; i++)
                for (int c = 0; c < nc; c++) jac[i * nc + c] *= scale[c];
        }
    };

    // IterationZero
    IterationLog it;
    it.iteration = 0;
    evaluate_gradient_and_jacobian(0);
    summary->initial_cost = x_cost;
    it.cost = x_cost;
    it.gradient_max_norm = gradient_max_norm;
    it.gradient_norm = gradient_norm;
    it.step_is_valid = true;
    it.step_is_successful = true;

    int num_consecutive_invalid_steps = 0;
    bool atleast_one_successful_step = false;

    // FinalizeIterationAndCheckIfMinimizerCanContinue
    auto finalize_and_can_continue = [&]() -> bool {
        it.trust_region_radius = radius;
        summary->iterations.push_back(it);
        if (it.iteration >= opt.max_num_iterations) { summary->termination = NO_CONVERGENCE; return false; }
        if (it.gradient_max_norm <= opt.gradient_tolerance) { summary->termination = CONV_GRADIENT; return false; }
        if (it.trust_region_radius <= opt.min_trust_region_radius) { summary->termination = CONV_RADIUS; return false; }
        return true;
    };

    while (finalize_and_can_continue()) {
        const double previous_gradient_norm = it.gradient_norm;
        const double previous_gradient_max_norm = it.gradient_max_norm;
        const int next_iteration = summary->iterations.back().iteration + 1;
        it = IterationLog();
        it.iteration = next_iteration;

        // ---- ComputeTrustRegionStep: LevenbergMarquardtStrategy::ComputeStep + DENSE_QR ----
        if (!reuse_diagonal) {
            for (int c = 0; c < nc; c++) {
                double s2 = 0.0;
                for (std::size_t i = 0; i < m; i++) s2 += jac[i * nc + c] * jac[i * nc + c];
                diagonal[c] = std::min(std::max(s2, opt.min_lm_diagonal), opt.max_lm_diagonal);
            }
        }
        double lm_diagonal[6];
        for (int c = 0; c < nc; c++) lm_diagonal[c] = std::sqrt(diagonal[c] / radius);
        std::vector<double> A((m + nc) * nc, 0.0), rhs(m + nc, 0.0);
        std::copy(jac.begin(), jac.end(), A.begin());
        std::copy(residuals.begin(), residuals.end(), rhs.begin());
        for (int c = 0; c < nc; c++) A[(m + c) * nc + c] = lm_diagonal[c];
        double step[6];
        const bool linear_ok = householder_solve6(A, rhs, m + nc, step);
        reuse_diagonal = true;
        double model_cost_change = 0.0;
        double delta[6] = {0, 0, 0, 0, 0, 0};
        if (linear_ok) {
            for (int c = 0; c < nc; c++) step[c] = -step[c];
            // model_cost_change = -(J s)'(r + J s / 2)
            for (std::size_t i = 0; i < m; i++) {
                double v = 0.0;
                for (int c = 0; c < nc; c++) v += jac[i * nc + c] * step[c];
                model_res[i] = v;
            }
            for (std::size_t i = 0; i < m; i++) model_cost_change += model_res[i] * (residuals[i] + model_res[i] / 2.0);
            model_cost_change = -model_cost_change;
            it.step_is_valid = model_cost_change > 0.0;
            if (it.step_is_valid) {
                for (int c = 0; c < nc; c++) delta[c] = step[c] * scale[c];
                num_consecutive_invalid_steps = 0;
            }
        }
        if (!it.step_is_valid) {  // HandleInvalidStep
            if (++num_consecutive_invalid_steps >= opt.max_num_consecutive_invalid_steps) {
                summary->termination = FAILURE;
                break;
            }
            radius = radius / decrease_factor;  // StepIsInvalid == StepRejected
            decrease_factor *= 2.0;
            reuse_diagonal = true;
            it.cost = x_cost;
            it.cost_change = 0.0;
            it.gradient_max_norm = previous_gradient_max_norm;
            it.gradient_norm = previous_gradient_norm;
            it.step_norm = 0.0;
            it.relative_decrease = 0.0;
            continue;
        }

        // ---- ComputeCandidatePointAndEvaluateCost ----
        double candidate_x[7], candidate_cost;
        Problem::plus(x, delta, candidate_x);
        problem.evaluate(candidate_x, &candidate_cost, nullptr, nullptr);
        summary->num_cost_evals++;
        if (!std::isfinite(candidate_cost)) candidate_cost = std::numeric_limits<double>::max();

        // ---- ParameterToleranceReached (computes step_norm as a side effect) ----
        if (!opt.param_tol_needs_successful_step || atleast_one_successful_step) {
            double sn = 0.0;
            for (int i = 0; i < 7; i++) sn += (x[i] - candidate_x[i]) * (x[i] - candidate_x[i]);
            it.step_norm = std::sqrt(sn);
            const double tol = opt.parameter_tolerance * (x_norm + opt.parameter_tolerance);
            if (it.step_norm <= tol) { summary->termination = CONV_PARAMETER; break; }
        }
        // ---- FunctionToleranceReached ----
        it.cost_change = x_cost - candidate_cost;
        if (std::fabs(it.cost_change) <= opt.function_tolerance * x_cost) { summary->termination = CONV_FUNCTION; break; }

        // ---- IsStepSuccessful (monotonic TrustRegionStepEvaluator) ----
        if (candidate_cost >= std::numeric_limits<double>::max())
            it.relative_decrease = std::numeric_limits<double>::lowest();
        else
            it.relative_decrease = (x_cost - candidate_cost) / model_cost_change;

        if (it.relative_decrease > opt.min_relative_decrease) {  // HandleSuccessfulStep
            for (int i = 0; i < 7; i++) x[i] = candidate_x[i];
            x_norm = 0.0;
            for (int i = 0; i < 7; i++) x_norm += x[i] * x[i];
            x_norm = std::sqrt(x_norm);
            evaluate_gradient_and_jacobian(it.iteration);
            it.cost = x_cost;
            it.gradient_max_norm = gradient_max_norm;
            it.gradient_norm = gradient_norm;
            it.step_is_successful = true;
            radius = radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * it.relative_decrease - 1.0, 3));
            radius = std::min(opt.max_trust_region_radius, radius);
            decrease_factor = 2.0;
            reuse_diagonal = false;
            atleast_one_successful_step = true;
        } else {
            it.step_is_successful = false;
            it.cost = candidate_cost;
            it.gradient_max_norm = previous_gradient_max_norm;
            it.gradient_norm = previous_gradient_norm;
            radius = radius / decrease_factor;
            decrease_factor *= 2.0;
            reuse_diagonal = true;
        }
    }

    summary->final_cost = x_cost;
    for (int i = 0; i < 7; i++) x_inout[i] = x[i];
}

}  // namespace mini_ceres

#endif  // LO_ORACLE_MINI_CERES_H
