// L-BFGS with a strong-Wolfe line search as a DEVICE-side state machine.
//
// The reference tracks every frame with torch.optim.LBFGS(history_size=5, line_search_fn='strong_wolfe')
// over 11 unknowns (gslam/frontend.py:604-662; the window poses likewise, gslam/backend.py:436-506): every
// comparison of the optimizer is a device->host sync.  gslam_b200/lbfgs.py restates the algorithm on the
// host (one sync per objective evaluation).  Here the optimizer itself lives on the GPU: its whole state
// is one struct in device memory and gsb_lbfgs_advance -- one thread, float64 -- consumes the (loss,
// gradient) of the point just evaluated and writes the next point to evaluate into `theta`.  The caller
// replays   [evaluate(theta) -> result ; gsb_lbfgs_advance]   max_eval times with NO host sync (once the
// search has finished, advance leaves theta at the final iterate and the replays are idle) and reads the
// outcome back once.  The arithmetic follows lbfgs.py statement by statement (which follows
// torch/optim/lbfgs.py); tests/test_gpu_lbfgs.py checks the iterates against it.
#include "common.cuh"

#define LB_MAX_N 96
#define LB_MAX_HIST 16

struct LbPoint {
    double t, f, gtd;
    double g[LB_MAX_N];
};

enum { LB_INIT = 0, LB_BRACKET = 1, LB_ZOOM = 2, LB_DONE = 3 };

struct LbState {
    int n, hist, max_iter, max_eval;
    double lr, tol_grad, tol_change;
    int phase, n_iter, evals, hist_len, it, max_ls, stalled, lo, hi, bracket_n, ls_done, ls_evals;
    double loss, prev_loss, t, gtd, h_diag, d_norm, t_req, first_loss, last_loss;
    double x[LB_MAX_N], x_init[LB_MAX_N], d[LB_MAX_N], grad[LB_MAX_N], prev_grad[LB_MAX_N];
    double ys[LB_MAX_HIST][LB_MAX_N], ss[LB_MAX_HIST][LB_MAX_N], rho[LB_MAX_HIST];
    LbPoint prev, nw, br[2], p0;
};

__device__ double lb_dot(const double* a, const double* b, int n) {
    double s = 0.0;
    for (int i = 0; i < n; ++i) s += a[i] * b[i];
    return s;
}
__device__ double lb_absmax(const double* a, int n) {
    double s = 0.0;
    for (int i = 0; i < n; ++i) s = fmax(s, fabs(a[i]));
    return s;
}
__device__ void lb_copy_point(LbPoint& dst, const LbPoint& src, int n) {
    dst.t = src.t; dst.f = src.f; dst.gtd = src.gtd;
    for (int i = 0; i < n; ++i) dst.g[i] = src.g[i];
}

// minimiser of the cubic through two points with slopes, clipped to [lo, hi]; the midpoint without a real one
__device__ double lb_cubic_min(double x1, double f1, double g1, double x2, double f2, double g2, double lo, double hi) {
    const double d1 = g1 + g2 - 3.0 * (f1 - f2) / (x1 - x2);
    const double disc = d1 * d1 - g1 * g2;
    if (disc < 0) return 0.5 * (lo + hi);
    const double d2 = sqrt(disc);
    double pos;
    if (x1 <= x2) pos = x2 - (x2 - x1) * ((g2 + d2 - d1) / (g2 - g1 + 2.0 * d2));
    else pos = x1 - (x1 - x2) * ((g1 + d2 - d1) / (g1 - g2 + 2.0 * d2));
    return fmin(fmax(pos, lo), hi);
}

__device__ void lb_request(LbState& S, float* theta, double t) {   // evaluate at x_init + t d next
    S.t_req = t;
    for (int i = 0; i < S.n; ++i) theta[i] = (float)(S.x_init[i] + t * S.d[i]);
}

__device__ void lb_finish(LbState& S, float* theta) {
    S.phase = LB_DONE;
    for (int i = 0; i < S.n; ++i) theta[i] = (float)S.x[i];
}

// head of the outer loop (lbfgs_minimize, `while n_iter < max_iter`) up to the first evaluation of its search
__device__ void lb_begin_iter(LbState& S, float* theta) {
    const int n = S.n;
    if (S.n_iter >= S.max_iter) { lb_finish(S, theta); return; }
    S.n_iter += 1;
    if (S.n_iter > 1) {
        double y[LB_MAX_N], s[LB_MAX_N];
        for (int i = 0; i < n; ++i) { y[i] = S.grad[i] - S.prev_grad[i]; s[i] = S.d[i] * S.t; }
        const double ys_dot = lb_dot(y, s, n);
        if (ys_dot > 1e-10) {
            if (S.hist_len == S.hist) {
                for (int h = 1; h < S.hist; ++h) {
                    for (int i = 0; i < n; ++i) { S.ys[h - 1][i] = S.ys[h][i]; S.ss[h - 1][i] = S.ss[h][i]; }
                    S.rho[h - 1] = S.rho[h];
                }
                S.hist_len -= 1;
            }
            const int h = S.hist_len++;
            for (int i = 0; i < n; ++i) { S.ys[h][i] = y[i]; S.ss[h][i] = s[i]; }
            S.rho[h] = 1.0 / ys_dot;
            S.h_diag = ys_dot / lb_dot(y, y, n);
        }
        double q[LB_MAX_N], alpha[LB_MAX_HIST];
        for (int i = 0; i < n; ++i) q[i] = -S.grad[i];
        for (int h = S.hist_len - 1; h >= 0; --h) {
            alpha[h] = lb_dot(S.ss[h], q, n) * S.rho[h];
            for (int i = 0; i < n; ++i) q[i] -= alpha[h] * S.ys[h][i];
        }
        for (int i = 0; i < n; ++i) S.d[i] = q[i] * S.h_diag;
        for (int h = 0; h < S.hist_len; ++h) {
            const double beta = lb_dot(S.ys[h], S.d, n) * S.rho[h];
            for (int i = 0; i < n; ++i) S.d[i] += (alpha[h] - beta) * S.ss[h][i];
        }
    }
    for (int i = 0; i < n; ++i) S.prev_grad[i] = S.grad[i];
    S.prev_loss = S.loss;
    double l1 = 0.0;
    for (int i = 0; i < n; ++i) l1 += fabs(S.grad[i]);
    S.t = S.n_iter == 1 ? fmin(1.0, 1.0 / l1) * S.lr : S.lr;
    S.gtd = lb_dot(S.grad, S.d, n);
    if (S.gtd > -S.tol_change) { lb_finish(S, theta); return; }
    for (int i = 0; i < n; ++i) S.x_init[i] = S.x[i];
    // strong_wolfe(...): first trial at t
    S.max_ls = S.max_eval - S.evals;
    S.d_norm = lb_absmax(S.d, n);
    S.p0.t = 0.0; S.p0.f = S.loss; S.p0.gtd = S.gtd;
    for (int i = 0; i < n; ++i) S.p0.g[i] = S.grad[i];
    lb_copy_point(S.prev, S.p0, n);
    S.it = 0; S.ls_evals = 0; S.ls_done = 0; S.bracket_n = 0; S.stalled = 0;
    S.phase = LB_BRACKET;
    lb_request(S, theta, S.t);
}

// the zoom loop of strong_wolfe up to its next evaluation; returns true when the search is over
__device__ bool lb_zoom_step(LbState& S, float* theta) {
    if (S.ls_done || S.it >= S.max_ls) return true;
    const LbPoint& a = S.br[0];
    const LbPoint& b = S.br[1];
    if (fabs(b.t - a.t) * S.d_norm < S.tol_change) return true;
    const double blo = fmin(a.t, b.t), bhi = fmax(a.t, b.t);
    double t_try = lb_cubic_min(a.t, a.f, a.gtd, b.t, b.f, b.gtd, blo, bhi);
    const double eps = 0.1 * (bhi - blo);
    if (fmin(bhi - t_try, t_try - blo) < eps) {
        if (S.stalled || t_try >= bhi || t_try <= blo) {
            t_try = fabs(t_try - bhi) < fabs(t_try - blo) ? bhi - eps : blo + eps;
            S.stalled = 0;
        } else {
            S.stalled = 1;
        }
    } else {
        S.stalled = 0;
    }
    S.phase = LB_ZOOM;
    lb_request(S, theta, t_try);
    return false;
}

// after the line search: accept its point and run the stop rules of the outer loop
__device__ void lb_after_search(LbState& S, float* theta) {
    const int n = S.n;
    const LbPoint& best = S.bracket_n > 1 ? S.br[S.lo] : S.br[0];
    S.loss = best.f; S.t = best.t;
    for (int i = 0; i < n; ++i) { S.grad[i] = best.g[i]; S.x[i] = S.x_init[i] + S.t * S.d[i]; }
    S.evals += S.ls_evals;
    if (S.n_iter == S.max_iter || S.evals >= S.max_eval) { lb_finish(S, theta); return; }
    if (lb_absmax(S.grad, n) <= S.tol_grad) { lb_finish(S, theta); return; }
    double step_max = 0.0;
    for (int i = 0; i < n; ++i) step_max = fmax(step_max, fabs(S.d[i] * S.t));
    if (step_max <= S.tol_change) { lb_finish(S, theta); return; }
    if (fabs(S.loss - S.prev_loss) < S.tol_change) { lb_finish(S, theta); return; }
    lb_begin_iter(S, theta);
}

__device__ void lb_enter_zoom(LbState& S, float* theta) {
    S.stalled = 0;
    if (S.bracket_n > 1) { S.lo = S.br[0].f <= S.br[1].f ? 0 : 1; S.hi = 1 - S.lo; }
    else { S.lo = 0; S.hi = 0; S.ls_done = 1; }
    if (lb_zoom_step(S, theta)) lb_after_search(S, theta);
}

__global__ void lbfgs_init_kernel(LbState* state, const float* __restrict__ theta, int n, int hist, int max_iter,
                                  int max_eval, double lr, double tol_grad, double tol_change) {
    LbState& S = *state;
    S.n = n; S.hist = hist; S.max_iter = max_iter; S.max_eval = max_eval;
    S.lr = lr; S.tol_grad = tol_grad; S.tol_change = tol_change;
    S.phase = LB_INIT; S.n_iter = 0; S.evals = 0; S.hist_len = 0; S.h_diag = 1.0; S.t = 0.0;
    S.first_loss = S.last_loss = 0.0;
    for (int i = 0; i < n; ++i) { S.x[i] = (double)theta[i]; S.x_init[i] = S.x[i]; S.d[i] = 0.0; }
}

// result = (loss, gradient[n]) of the point in theta; theta <- the next point to evaluate (or the final iterate)
__global__ void lbfgs_advance_kernel(LbState* state, float* __restrict__ theta, const float* __restrict__ result,
                                     float* __restrict__ summary) {
    LbState& S = *state;
    const int n = S.n;
    if (S.phase != LB_DONE) {
        const double f = (double)result[0];
        S.last_loss = f;
        if (S.phase == LB_INIT) {
            S.loss = f; S.first_loss = f; S.evals = 1;
            for (int i = 0; i < n; ++i) { S.grad[i] = (double)result[1 + i]; S.d[i] = -S.grad[i]; S.prev_grad[i] = S.grad[i]; }
            if (lb_absmax(S.grad, n) <= S.tol_grad) lb_finish(S, theta);
            else lb_begin_iter(S, theta);
        } else {
            // a trial point of the line search came back
            LbPoint& nw = S.nw;
            nw.t = S.t_req; nw.f = f;
            for (int i = 0; i < n; ++i) nw.g[i] = (double)result[1 + i];
            nw.gtd = lb_dot(nw.g, S.d, n);
            S.ls_evals += 1;
            if (S.phase == LB_BRACKET) {
                const double f0 = S.p0.f, gtd0 = S.p0.gtd;
                bool to_zoom = true;
                if (S.it < S.max_ls) {
                    if (nw.f > f0 + 1e-4 * nw.t * gtd0 || (S.it > 1 && nw.f >= S.prev.f)) {
                        lb_copy_point(S.br[0], S.prev, n); lb_copy_point(S.br[1], nw, n); S.bracket_n = 2;
                    } else if (fabs(nw.gtd) <= -0.9 * gtd0) {
                        lb_copy_point(S.br[0], nw, n); S.bracket_n = 1; S.ls_done = 1;
                    } else if (nw.gtd >= 0) {
                        lb_copy_point(S.br[0], S.prev, n); lb_copy_point(S.br[1], nw, n); S.bracket_n = 2;
                    } else {
                        // extrapolate: the next trial lies in [t + 0.01 (t - t_prev), 10 t]
                        const double t_next = lb_cubic_min(S.prev.t, S.prev.f, S.prev.gtd, nw.t, nw.f, nw.gtd,
                                                           nw.t + 0.01 * (nw.t - S.prev.t), nw.t * 10.0);
                        lb_copy_point(S.prev, nw, n);
                        S.it += 1;
                        lb_request(S, theta, t_next);
                        to_zoom = false;
                    }
                } else {
                    lb_copy_point(S.br[0], S.p0, n); lb_copy_point(S.br[1], nw, n); S.bracket_n = 2;
                }
                if (to_zoom) lb_enter_zoom(S, theta);
            } else {    // LB_ZOOM
                S.it += 1;
                const double f0 = S.p0.f, gtd0 = S.p0.gtd;
                if (nw.f > f0 + 1e-4 * nw.t * gtd0 || nw.f >= S.br[S.lo].f) {
                    lb_copy_point(S.br[S.hi], nw, n);
                    S.lo = S.br[0].f <= S.br[1].f ? 0 : 1; S.hi = 1 - S.lo;
                } else {
                    if (fabs(nw.gtd) <= -0.9 * gtd0) S.ls_done = 1;
                    else if (nw.gtd * (S.br[S.hi].t - S.br[S.lo].t) >= 0) lb_copy_point(S.br[S.hi], S.br[S.lo], n);
                    lb_copy_point(S.br[S.lo], nw, n);
                }
                if (lb_zoom_step(S, theta)) lb_after_search(S, theta);
            }
        }
    }
    if (summary) {      // what the host reads once at the end: done, evaluations, iterations, first / last loss
        summary[0] = S.phase == LB_DONE ? 1.f : 0.f;
        summary[1] = (float)(S.evals + (S.phase == LB_DONE ? 0 : S.ls_evals));
        summary[2] = (float)S.n_iter;
        summary[3] = (float)S.first_loss;
        summary[4] = (float)S.last_loss;
    }
}

GSB_API size_t gsb_lbfgs_state_bytes(void) { return sizeof(LbState); }

GSB_API int gsb_lbfgs_init(void* state, const float* theta, int n, int history_size, int max_iter, int max_eval,
                           double lr, double tolerance_grad, double tolerance_change, void* stream) {
    GSB_CHECK_ARG(state && theta && n >= 1 && n <= LB_MAX_N && history_size >= 1 && history_size <= LB_MAX_HIST);
    GSB_CHECK_ARG(max_iter >= 1 && max_eval >= 1);
    lbfgs_init_kernel<<<1, 1, 0, gsb_stream(stream)>>>(reinterpret_cast<LbState*>(state), theta, n, history_size,
                                                      max_iter, max_eval, lr, tolerance_grad, tolerance_change);
    GSB_LAUNCH_CHECK();
    return 0;
}

GSB_API int gsb_lbfgs_advance(void* state, float* theta, const float* result, float* summary, void* stream) {
    GSB_CHECK_ARG(state && theta && result);
    lbfgs_advance_kernel<<<1, 1, 0, gsb_stream(stream)>>>(reinterpret_cast<LbState*>(state), theta, result, summary);
    GSB_LAUNCH_CHECK();
    return 0;
}
