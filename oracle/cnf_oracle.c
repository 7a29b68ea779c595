/* cnf_oracle.c — CPU restatement (FP32, OpenMP over sample tiles) of the reference's
 * batched augmented-dynamics path.  TEST INFRASTRUCTURE ONLY: it may be imported/linked
 * by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs,
 * never by the product path (continuousnf.jl_b200/).
 *
 * PARITY UNPINNED: the reference (Julia, v0.26.0) cannot run in this environment and its
 * tests hold no golden vector for this path.  Functions cite
 * the reference lines they restate; the solver / Dense / AD semantics live in un-vendored
 * third-party packages (OrdinaryDiffEqDefault "1", Lux "1", NNlib "0.9",
 * DifferentiationInterface "0.6", Enzyme "0.13", SciMLSensitivity "7": Project.toml:41-70)
 * and follow their published algorithms as written down in SURVEY.md Appendix A.
 *
 * Build: gcc -O3 -march=native -fopenmp -shared -fPIC (see oracle/Makefile).
 */
#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "../include/cnfb200.h"
#include "../include/cnf_tsit5.h"

#define TS 16 /* samples per tile: inner loops vectorise over it */

/* NNlib.tanh_fast for Float32; Lux uses it for Dense(tanh) on the CPU */
static int g_perturb = 0;
void cnf_oracle_set_perturb(int on) { g_perturb = on; }

static inline float tanh_fast(float x) {
    if (g_perturb) return tanhf(x);
    float x2 = x * x;
    float n = fmaf(x2, fmaf(x2, fmaf(x2, fmaf(x2, CNF_TANH_N4, CNF_TANH_N3), CNF_TANH_N2), CNF_TANH_N1), 1.0f);
    float d = fmaf(x2, fmaf(x2, fmaf(x2, fmaf(x2, CNF_TANH_D4, CNF_TANH_D3), CNF_TANH_D2), CNF_TANH_D1), 1.0f);
    if (x2 < CNF_TANH_CLAMP) return x * (n / d);
    return x > 0.0f ? 1.0f : -1.0f;
}

typedef struct {
    cnf_config cfg;
    const float* theta;
    int64_t w_off[CNF_MAX_LAYERS], b_off[CNF_MAX_LAYERS];
    int D, C, L, maxd;
} omodel;

static void omodel_init(omodel* m, const cnf_config* cfg, const float* theta,
                        const int64_t* w_off, const int64_t* b_off) {
    m->cfg = *cfg;
    m->theta = theta;
    m->D = cfg->D;
    m->C = cfg->C;
    m->L = cfg->n_layers;
    m->maxd = 0;
    for (int l = 0; l <= m->L; ++l)
        if (cfg->dims[l] > m->maxd) m->maxd = cfg->dims[l];
    for (int l = 0; l < m->L; ++l) {
        m->w_off[l] = w_off[l];
        m->b_off[l] = b_off[l];
    }
}

/* per-thread scratch: activations h_0..h_L (SoA d x TS) and working vectors */
typedef struct {
    float* h[CNF_MAX_LAYERS + 1];
    float* g[CNF_MAX_LAYERS + 1]; /* VJP cotangents g_l / JVP tangents tau_l per layer */
    float* q[CNF_MAX_LAYERS + 1]; /* q_l = W_l^T g_l (VJP) or r_l = W_l tau_{l-1} (JVP) */
    float* t0;
    float* t1;
    float* t2;
    float* zero; /* stands in for h_L where its derivative 1 - h^2 is meant and the last layer is linear */
    float* store;
} oscratch;

static void oscratch_alloc(oscratch* s, const omodel* m) {
    size_t per = (size_t)m->maxd * TS;
    size_t n = (size_t)(3 * (CNF_MAX_LAYERS + 1) + 4) * per;
    s->store = (float*)aligned_alloc(64, n * sizeof(float));
    float* p = s->store;
    for (int l = 0; l <= CNF_MAX_LAYERS; ++l) { s->h[l] = p; p += per; }
    for (int l = 0; l <= CNF_MAX_LAYERS; ++l) { s->g[l] = p; p += per; }
    for (int l = 0; l <= CNF_MAX_LAYERS; ++l) { s->q[l] = p; p += per; }
    s->t0 = p; p += per;
    s->t1 = p; p += per;
    s->t2 = p; p += per;
    s->zero = p;
    memset(s->zero, 0, per * sizeof(float));
}
/* h_l as the argument of the activation derivative: d_l = 1 - h_l^2, or 1 for a linear last layer
 * (CNF_FLAG_LAST_LINEAR: PlanarLayer's u * act(.), src/layers/planar_layer.jl:70-88) */
#define H_DER(m, sc, l) ((((m)->cfg.flags & CNF_FLAG_LAST_LINEAR) && (l) == (m)->L) ? (const float*)(sc)->zero : (const float*)(sc)->h[l])
static void oscratch_free(oscratch* s) { free(s->store); }

/* out[o][:] = (bias ? b[o] : 0) + sum_{i<din_used} W[o,i] * in[i][:]   (W: dout x din col-major) */
static void dense_mul(const float* W, const float* b, int dout, int din_used,
                      const float* in, float* out) {
    /* accumulation order per output: bias, then ascending i (same in both branches) */
    if ((long)dout * din_used <= 65536) { /* small layer: accumulators stay in registers */
        for (int o = 0; o < dout; ++o) {
            float acc[TS];
            float bv = b ? b[o] : 0.0f;
            for (int s = 0; s < TS; ++s) acc[s] = bv;
            for (int i = 0; i < din_used; ++i) {
                float w = W[(size_t)i * dout + o];
                const float* x = in + i * TS;
                for (int s = 0; s < TS; ++s) acc[s] = fmaf(w, x[s], acc[s]);
            }
            for (int s = 0; s < TS; ++s) out[o * TS + s] = acc[s];
        }
        return;
    }
    for (int o = 0; o < dout; ++o) {
        float bv = b ? b[o] : 0.0f;
        for (int s = 0; s < TS; ++s) out[o * TS + s] = bv;
    }
    for (int i = 0; i < din_used; ++i) {
        const float* wc = W + (size_t)i * dout;
        const float* x = in + i * TS;
        for (int o = 0; o < dout; ++o) {
            float w = wc[o];
            float* y = out + o * TS;
            for (int s = 0; s < TS; ++s) y[s] = fmaf(w, x[s], y[s]);
        }
    }
}

/* out[i][:] = sum_o W[o,i] * in[o][:]  for i < din_used */
static void dense_mulT(const float* W, int dout, int din_used, const float* in, float* out) {
    for (int i = 0; i < din_used; ++i) {
        const float* wc = W + (size_t)i * dout;
        float acc[TS];
        for (int s = 0; s < TS; ++s) acc[s] = 0.0f;
        for (int o = 0; o < dout; ++o) {
            float w = wc[o];
            const float* x = in + o * TS;
            for (int s = 0; s < TS; ++s) acc[s] = fmaf(w, x[s], acc[s]);
        }
        for (int s = 0; s < TS; ++s) out[i * TS + s] = acc[s];
    }
}

/* forward chain h_l = tanh_fast(W_l h_{l-1} + b_l) with h_0 = [z; ys] already in sc->h[0]
 * (Lux Chain(Dense(.,tanh)...), CondLayer: src/layers/cond_layer.jl:1-9) */
static void nn_forward(const omodel* m, oscratch* sc) {
    for (int l = 0; l < m->L; ++l) {
        int din = m->cfg.dims[l], dout = m->cfg.dims[l + 1];
        const int lin = (m->cfg.flags & CNF_FLAG_LAST_LINEAR) && l == m->L - 1;
        dense_mul(m->theta + m->w_off[l], lin ? NULL : m->theta + m->b_off[l], dout, din, sc->h[l], sc->h[l + 1]);
        float* y = sc->h[l + 1];
        if (!lin)
            for (int k = 0; k < dout * TS; ++k) y[k] = tanh_fast(y[k]);
    }
}

/* reverse chain (VecJac): g_L = seed .* (1-h_L^2); q_l = W_l^T g_l; g_{l-1} = q_l .* (1-h_{l-1}^2).
 * Result eps^T J = q_1[0:D].  (Enzyme pullback of the Dense chain, src/icnf.jl:331-332; A.6) */
static void vjp_chain(const omodel* m, oscratch* sc, const float* seed) {
    int L = m->L;
    {
        int d = m->cfg.dims[L];
        const float* h = H_DER(m, sc, L);
        float* g = sc->g[L];
        for (int k = 0; k < d * TS; ++k) g[k] = seed[k] * fmaf(-h[k], h[k], 1.0f);
    }
    for (int l = L; l >= 1; --l) {
        int din = m->cfg.dims[l - 1], dout = m->cfg.dims[l];
        int used = (l == 1) ? m->D : din;
        dense_mulT(m->theta + m->w_off[l - 1], dout, used, sc->g[l], sc->q[l]);
        if (l > 1) {
            const float* h = sc->h[l - 1];
            const float* q = sc->q[l];
            float* g = sc->g[l - 1];
            for (int k = 0; k < din * TS; ++k) g[k] = q[k] * fmaf(-h[k], h[k], 1.0f);
        }
    }
}

/* forward chain (JacVec): r_1 = W_1[:,0:D] seed; tau_l = r_l .* (1-h_l^2); r_{l+1} = W_{l+1} tau_l.
 * Result J eps = tau_L = g[L].  (pushforward, src/icnf.jl:397-402; A.6) */
static void jvp_chain(const omodel* m, oscratch* sc, const float* seed) {
    const float* in = seed;
    for (int l = 1; l <= m->L; ++l) {
        int din = m->cfg.dims[l - 1], dout = m->cfg.dims[l];
        int used = (l == 1) ? m->D : din;
        dense_mul(m->theta + m->w_off[l - 1], NULL, dout, used, in, sc->q[l]);
        const float* h = H_DER(m, sc, l);
        const float* r = sc->q[l];
        float* t = sc->g[l];
        for (int k = 0; k < dout * TS; ++k) t[k] = r[k] * fmaf(-h[k], h[k], 1.0f);
        in = sc->g[l];
    }
}

static void load_h0(const omodel* m, oscratch* sc, const float* u, int R, const float* ys,
                    int64_t b0, int ns) {
    int D = m->D, C = m->C;
    float* h0 = sc->h[0];
    for (int i = 0; i < D; ++i)
        for (int s = 0; s < TS; ++s) h0[i * TS + s] = (s < ns) ? u[(b0 + s) * R + i] : 0.0f;
    for (int i = 0; i < C; ++i)
        for (int s = 0; s < TS; ++s) h0[(D + i) * TS + s] = (s < ns) ? ys[(b0 + s) * C + i] : 0.0f;
}

/* augmented_f on one tile of samples.
 * TestMode:  du = [zdot; -tr J]               (src/icnf.jl:148-164, src/utils.jl:1-36)
 * TrainMode: du = [zdot; -eps.v; ||zdot||; ||v||], v = J^T eps (VecJac, src/icnf.jl:318-350)
 *                                               or v = J eps   (JacVec, src/icnf.jl:384-420) */
static void rhs_tile(const omodel* m, oscratch* sc, int mode, const float* u, const float* ys,
                     const float* eps, float* du, int64_t b0, int ns) {
    int D = m->D;
    int R = D + (mode == CNF_MODE_TRAIN ? 3 : 1);
    load_h0(m, sc, u, R, ys, b0, ns);
    nn_forward(m, sc);
    const float* y = sc->h[m->L];
    float ldot[TS], nrmv[TS];
    for (int s = 0; s < TS; ++s) ldot[s] = 0.0f, nrmv[s] = 0.0f;
    if (mode == CNF_MODE_TEST) {
        /* jacobian_batched: D one-hot sweeps; only the diagonal enters -tr(J) */
        for (int i = 0; i < D; ++i) {
            float* seed = sc->t0;
            memset(seed, 0, sizeof(float) * (size_t)D * TS);
            for (int s = 0; s < TS; ++s) seed[i * TS + s] = 1.0f;
            if (m->cfg.trace_mode == CNF_TRACE_VJP) {
                vjp_chain(m, sc, seed);
                for (int s = 0; s < TS; ++s) ldot[s] -= sc->q[1][i * TS + s];
            } else {
                jvp_chain(m, sc, seed);
                for (int s = 0; s < TS; ++s) ldot[s] -= sc->g[m->L][i * TS + s];
            }
        }
    } else {
        float* e = sc->t0;
        for (int i = 0; i < D; ++i)
            for (int s = 0; s < TS; ++s) e[i * TS + s] = (s < ns) ? eps[(b0 + s) * D + i] : 0.0f;
        const float* v;
        if (m->cfg.trace_mode == CNF_TRACE_VJP) {
            vjp_chain(m, sc, e);
            v = sc->q[1];
        } else {
            jvp_chain(m, sc, e);
            v = sc->g[m->L];
        }
        for (int i = 0; i < D; ++i)
            for (int s = 0; s < TS; ++s) {
                ldot[s] = fmaf(-v[i * TS + s], e[i * TS + s], ldot[s]);
                nrmv[s] = fmaf(v[i * TS + s], v[i * TS + s], nrmv[s]);
            }
    }
    for (int s = 0; s < ns; ++s) {
        float* o = du + (b0 + s) * R;
        float ny = 0.0f;
        for (int i = 0; i < D; ++i) {
            float yi = y[i * TS + s];
            o[i] = yi;
            ny = fmaf(yi, yi, ny);
        }
        o[D] = ldot[s];
        if (mode == CNF_MODE_TRAIN) {
            o[D + 1] = (m->cfg.flags & CNF_FLAG_NORM_Z) ? sqrtf(ny) : 0.0f;
            o[D + 2] = (m->cfg.flags & CNF_FLAG_NORM_J) ? sqrtf(nrmv[s]) : 0.0f;
        }
    }
}

/* du = augmented_f(u) over the whole batch */
static void rhs_all(const omodel* m, int mode, const float* u, const float* ys, const float* eps,
                    float* du, int64_t B) {
    int64_t ntile = (B + TS - 1) / TS;
#pragma omp parallel
    {
        oscratch sc;
        oscratch_alloc(&sc, m);
#pragma omp for schedule(static)
        for (int64_t t = 0; t < ntile; ++t) {
            int64_t b0 = t * TS;
            int ns = (int)((B - b0) < TS ? (B - b0) : TS);
            rhs_tile(m, &sc, mode, u, ys, eps, du, b0, ns);
        }
        oscratch_free(&sc);
    }
}

int32_t cnf_oracle_augmented_f(const cnf_config* cfg, const float* theta, const int64_t* w_off,
                               const int64_t* b_off, int32_t mode, const float* u,
                               const float* ys, const float* eps, int64_t B, float* du) {
    omodel m;
    omodel_init(&m, cfg, theta, w_off, b_off);
    rhs_all(&m, mode, u, ys, eps, du, B);
    return 0;
}

/* ------------------------------------------------------------------------------------
 * Adaptive Tsit5 5(4) FSAL solve over the matrix-valued state;
 * call site src/base_icnf.jl:137-143.  State u: R x B sample-major.  One error norm for
 * the whole matrix (A.3).  Controller scalars are evaluated in double and stored as float.
 * ------------------------------------------------------------------------------------ */
typedef struct {
    int64_t naccept, nreject, nf;
    int32_t retcode;
    float t_final, dt_last;
    float stiff_max;   /* largest stiffness estimate of the default algorithm's auto-switch (reported only) */
    int32_t stiff_run; /* longest run of consecutive accepted steps above the switch tolerance */
} ostats;

typedef struct { /* accepted-step tape for the discrete adjoint */
    float** u;   /* u_n (R x B) at the start of accepted step n */
    float* dt;
    int64_t n, cap;
} otape;

static void tape_push(otape* tp, const float* u, size_t nel, float dt) {
    if (tp->n == tp->cap) {
        tp->cap = tp->cap ? 2 * tp->cap : 64;
        tp->u = (float**)realloc(tp->u, sizeof(float*) * (size_t)tp->cap);
        tp->dt = (float*)realloc(tp->dt, sizeof(float) * (size_t)tp->cap);
    }
    tp->u[tp->n] = (float*)malloc(nel * sizeof(float));
    memcpy(tp->u[tp->n], u, nel * sizeof(float));
    tp->dt[tp->n] = dt;
    tp->n++;
}
static void tape_free(otape* tp) {
    for (int64_t i = 0; i < tp->n; ++i) free(tp->u[i]);
    free(tp->u);
    free(tp->dt);
    memset(tp, 0, sizeof(*tp));
}

static const float A_[7][6] = {
    {0},
    {(float)CNF_A21},
    {(float)CNF_A31, (float)CNF_A32},
    {(float)CNF_A41, (float)CNF_A42, (float)CNF_A43},
    {(float)CNF_A51, (float)CNF_A52, (float)CNF_A53, (float)CNF_A54},
    {(float)CNF_A61, (float)CNF_A62, (float)CNF_A63, (float)CNF_A64, (float)CNF_A65},
    {(float)CNF_A71, (float)CNF_A72, (float)CNF_A73, (float)CNF_A74, (float)CNF_A75, (float)CNF_A76}};
static const float BT_[7] = {(float)CNF_BT1, (float)CNF_BT2, (float)CNF_BT3, (float)CNF_BT4,
                             (float)CNF_BT5, (float)CNF_BT6, (float)CNF_BT7};

/* s = u + dt * sum_{j<nk} a[j] k[j]   (stage input / new state) */
static void rk_combine(float* s, const float* u, float* const* k, const float* a, int nk,
                       float dt, int64_t nel) {
#pragma omp parallel for schedule(static)
    for (int64_t e = 0; e < nel; ++e) {
        float acc = a[0] * k[0][e];
        for (int j = 1; j < nk; ++j) acc = fmaf(a[j], k[j][e], acc);
        s[e] = fmaf(dt, acc, u[e]);
    }
}

static double sumsq_scaled(const float* x, const float* y, const float* a, const float* b,
                           float abstol, float reltol, int64_t nel) {
    /* sum ((x - y) / (abstol + max(|a|,|b|) reltol))^2 ; y, b may be NULL */
    double acc = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : acc)
    for (int64_t e = 0; e < nel; ++e) {
        float v = y ? x[e] - y[e] : x[e];
        float m = fabsf(a[e]);
        if (b) m = fmaxf(m, fabsf(b[e]));
        float r = v / fmaf(m, reltol, abstol);
        acc += (double)r * (double)r;
    }
    return acc;
}

static int has_nan(const float* x, int64_t nel) {
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad)
    for (int64_t e = 0; e < nel; ++e) bad |= (x[e] != x[e]);
    return bad;
}

/* accepted dt sequence of the most recent solve (for the frozen-step f64 gradient check) */
static float g_last_dts[4096];
static int64_t g_last_ndts = 0;
int64_t cnf_oracle_last_dts(float* out, int64_t cap) {
    int64_t n = g_last_ndts < cap ? g_last_ndts : cap;
    for (int64_t i = 0; i < n; ++i) out[i] = g_last_dts[i];
    return g_last_ndts;
}

static void tsit5_solve(const omodel* m, int mode, float* u, const float* ys, const float* eps,
                        int64_t B, float t0, float t1, ostats* st, otape* tape) {
    const int R = m->D + (mode == CNF_MODE_TRAIN ? 3 : 1);
    const int64_t nel = (int64_t)R * B;
    const float reltol = m->cfg.reltol > 0 ? m->cfg.reltol : CNF_DEFAULT_RELTOL;
    const float abstol = m->cfg.abstol > 0 ? m->cfg.abstol : CNF_DEFAULT_ABSTOL;
    const int64_t maxiters = m->cfg.maxiters > 0 ? m->cfg.maxiters : CNF_DEFAULT_MAXITERS;
    const float tdir = (t1 >= t0) ? 1.0f : -1.0f;
    const float dtmax = m->cfg.dtmax > 0 ? m->cfg.dtmax : fabsf(t1 - t0);
    float* k[7];
    for (int j = 0; j < 7; ++j) k[j] = (float*)malloc(sizeof(float) * (size_t)nel);
    float* s = (float*)malloc(sizeof(float) * (size_t)nel);
    float* unew = (float*)malloc(sizeof(float) * (size_t)nel);
    memset(st, 0, sizeof(*st));

    rhs_all(m, mode, u, ys, eps, k[0], B);
    st->nf = 1;
    float dt;
    if (m->cfg.dt0 > 0) {
        dt = tdir * m->cfg.dt0;
    } else { /* Hairer-Wanner initial step, A.5 */
        double d0 = sqrt(sumsq_scaled(u, NULL, u, NULL, abstol, reltol, nel) / (double)nel);
        double d1 = sqrt(sumsq_scaled(k[0], NULL, u, NULL, abstol, reltol, nel) / (double)nel);
        double dta = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
        if (dta > dtmax) dta = dtmax;
        float dt0f = (float)dta;
        float one = 1.0f;
        float* kk[1] = {k[0]};
        rk_combine(s, u, kk, &one, 1, tdir * dt0f, nel);
        rhs_all(m, mode, s, ys, eps, k[1], B);
        double d2 = sqrt(sumsq_scaled(k[1], k[0], u, NULL, abstol, reltol, nel) / (double)nel) / (double)dt0f;
        double mx = d1 > d2 ? d1 : d2;
        double dt1 = (mx <= 1e-15) ? fmax(1e-6, (double)dt0f * 1e-3) : pow(10.0, -(2.0 + log10(mx)) / 5.0);
        double dtb = fmin(fmin(100.0 * (double)dt0f, dt1), (double)dtmax);
        dt = tdir * (float)dtb;
        st->nf += 2; /* f(u1) and the integrator's own re-evaluation of f(u0) (reused here) */
    }

    g_last_ndts = 0;
    float t = t0;
    int32_t stiff_cur = 0;
    st->stiff_max = -1.0f;
    double qold = CNF_QOLDINIT;
    int64_t iter = 0;
    st->retcode = CNF_OK;
    while (t != t1) {
        if (++iter > maxiters) { st->retcode = CNF_RET_MAXITERS; break; }
        float adt = fminf(fabsf(dt), dtmax);
        adt = fminf(adt, fabsf(t1 - t));
        dt = tdir * adt;
        if (t + dt == t || (m->cfg.dtmin > 0 && adt < m->cfg.dtmin)) { st->retcode = CNF_RET_DTLESSTHANMIN; break; }
        for (int i = 1; i < 6; ++i) {
            rk_combine(s, u, k, A_[i], i, dt, nel);
            rhs_all(m, mode, s, ys, eps, k[i], B);
        }
        rk_combine(unew, u, k, A_[6], 6, dt, nel);
        rhs_all(m, mode, unew, ys, eps, k[6], B);
        st->nf += 6;
        /* utilde = dt * sum btilde_j k_j ; EEst = RMS(utilde / (abstol + max(|u|,|unew|) reltol)) */
        double acc = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : acc)
        for (int64_t e = 0; e < nel; ++e) {
            float a = BT_[0] * k[0][e];
            for (int j = 1; j < 7; ++j) a = fmaf(BT_[j], k[j][e], a);
            float ut = dt * a;
            float sc = fmaf(fmaxf(fabsf(u[e]), fabsf(unew[e])), reltol, abstol);
            float r = ut / sc;
            acc += (double)r * (double)r;
        }
        float EEst = (float)sqrt(acc / (double)nel);
        if (EEst != EEst || has_nan(unew, nel)) { st->retcode = CNF_RET_UNSTABLE; break; }
        if (getenv("CNF_ORACLE_DEBUG")) fprintf(stderr, "iter %lld t %.7g dt %.7g EEst %.7g qold %.4g\n", (long long)iter, t, dt, EEst, qold);
        double q11 = 0.0, q;
        if (EEst == 0.0f) {
            q = 1.0 / CNF_QMAX;
        } else {
            q11 = pow((double)EEst, CNF_BETA1);
            q = q11 / pow(qold, CNF_BETA2);
            q = fmax(1.0 / CNF_QMAX, fmin(1.0 / CNF_QMIN, q / CNF_GAMMA));
        }
        if (EEst <= 1.0f) {
            /* A.1, reported only: OrdinaryDiffEq's composite default would test |dt| * eigen_est / 3.5068 > 9/10 with
             * eigen_est = ||k7 - k6|| / ||u_new - g6|| (RMS norms; g6 = the 6th stage input) after every step */
            {
                double su = 0.0, sd = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : su, sd)
                for (int64_t e = 0; e < nel; ++e) {
                    float gd = (A_[6][0] - A_[5][0]) * k[0][e];
                    for (int j = 1; j < 5; ++j) gd = fmaf(A_[6][j] - A_[5][j], k[j][e], gd);
                    gd = dt * fmaf(A_[6][5], k[5][e], gd);
                    float kd = k[6][e] - k[5][e];
                    su += (double)kd * (double)kd;
                    sd += (double)gd * (double)gd;
                }
                if (sd > 0.0) {
                    float stiff = (float)(fabs((double)dt) * sqrt(su / sd) / CNF_TSIT5_STABILITY);
                    if (stiff > st->stiff_max) st->stiff_max = stiff;
                    stiff_cur = stiff > CNF_STIFFTOL ? stiff_cur + 1 : 0;
                    if (stiff_cur > st->stiff_run) st->stiff_run = stiff_cur;
                }
            }
            if (tape) tape_push(tape, u, (size_t)nel, dt);
            if (st->naccept < 4096) g_last_dts[st->naccept] = dt;
            g_last_ndts = st->naccept + 1;
            float tn = t + dt;
            if (fabsf(tn - t1) < 100.0f * 1.1920929e-7f * fmaxf(fabsf(t), fabsf(t1))) tn = t1;
            t = tn;
            memcpy(u, unew, sizeof(float) * (size_t)nel);
            float* kt = k[0]; k[0] = k[6]; k[6] = kt;
            qold = fmax((double)EEst, CNF_QOLDINIT);
            st->dt_last = dt;
            dt = (float)((double)dt / q);
            st->naccept++;
        } else {
            dt = (float)((double)dt / fmin(1.0 / CNF_QMIN, q11 / CNF_GAMMA));
            st->nreject++;
        }
    }
    st->t_final = t;
    for (int j = 0; j < 7; ++j) free(k[j]);
    free(s);
    free(unew);
}

/* u0 = [xs; 0]  (inference_prob, src/base_icnf.jl:266-286 / :288-309) */
static float* assemble_u0(const omodel* m, int R, const float* xs, int64_t B) {
    float* u = (float*)calloc((size_t)R * (size_t)B, sizeof(float));
    int nv = m->cfg.nvars;
    for (int64_t b = 0; b < B; ++b)
        for (int i = 0; i < nv; ++i) u[b * R + i] = xs[b * nv + i];
    return u;
}

/* inference_sol (src/base_icnf.jl:167-189): split the final state, standard-normal logpdf,
 * logp = logpz - dlogp, A = ||z_aug|| if lambda3 != 0 and augmented.
 * regs rows: TrainMode (E, n, A) 3 x B; TestMode (A) 1 x B. */
static void final_split(const omodel* m, int mode, const float* u, int64_t B, float* logp,
                        float* regs, float* z_out) {
    int D = m->D, nv = m->cfg.nvars;
    int R = D + (mode == CNF_MODE_TRAIN ? 3 : 1);
    int nreg = mode == CNF_MODE_TRAIN ? 3 : 1;
    int useA = (m->cfg.flags & CNF_FLAG_NORM_Z_AUG) && (m->cfg.flags & CNF_FLAG_AUGMENTED);
    const float log2pi = 1.8378770664093453f;
    for (int64_t b = 0; b < B; ++b) {
        const float* ub = u + b * R;
        float zz = 0.0f, za = 0.0f;
        for (int i = 0; i < D; ++i) {
            zz = fmaf(ub[i], ub[i], zz);
            if (i >= nv) za = fmaf(ub[i], ub[i], za);
            if (z_out) z_out[b * D + i] = ub[i];
        }
        float logpz = -0.5f * zz - 0.5f * (float)D * log2pi;
        logp[b] = logpz - ub[D];
        if (regs) {
            float A = useA ? sqrtf(za) : 0.0f;
            if (mode == CNF_MODE_TRAIN) {
                regs[b * nreg + 0] = ub[D + 1];
                regs[b * nreg + 1] = ub[D + 2];
                regs[b * nreg + 2] = A;
            } else {
                regs[b * nreg] = A;
            }
        }
    }
}

static void copy_stats(cnf_stats* out, const ostats* st) {
    if (!out) return;
    memset(out, 0, sizeof(*out));
    out->naccept = st->naccept;
    out->nreject = st->nreject;
    out->nf = st->nf;
    out->retcode = st->retcode;
    out->t_final = st->t_final;
    out->dt_last = st->dt_last;
    out->stiff_max = st->stiff_max;
    out->stiff_run = st->stiff_run;
}

/* inference(icnf, mode, xs[, ys], ps, st)  (src/base_icnf.jl:407-426) */
int32_t cnf_oracle_logdensity(const cnf_config* cfg, const float* theta, const int64_t* w_off,
                              const int64_t* b_off, int32_t mode, const float* xs,
                              const float* ys, const float* eps, int64_t B, float t0, float t1,
                              float* logp, float* regs, float* z_out, cnf_stats* stats) {
    omodel m;
    omodel_init(&m, cfg, theta, w_off, b_off);
    int R = m.D + (mode == CNF_MODE_TRAIN ? 3 : 1);
    float* u = assemble_u0(&m, R, xs, B);
    ostats st;
    tsit5_solve(&m, mode, u, ys, eps, B, t0, t1, &st, NULL);
    final_split(&m, mode, u, B, logp, regs, z_out);
    copy_stats(stats, &st);
    free(u);
    return st.retcode;
}

/* generate(...) (src/base_icnf.jl:358-405, 202-211): solve from t_from to t_to starting at
 * [z0; 0], return rows 1:nvars */
int32_t cnf_oracle_generate(const cnf_config* cfg, const float* theta, const int64_t* w_off,
                            const int64_t* b_off, int32_t mode, const float* z0, const float* ys,
                            const float* eps, int64_t n, float t_from, float t_to,
                            float* xs_out, cnf_stats* stats) {
    omodel m;
    omodel_init(&m, cfg, theta, w_off, b_off);
    int D = m.D, nv = cfg->nvars;
    int R = D + (mode == CNF_MODE_TRAIN ? 3 : 1);
    float* u = (float*)calloc((size_t)R * (size_t)n, sizeof(float));
    for (int64_t b = 0; b < n; ++b)
        for (int i = 0; i < D; ++i) u[b * R + i] = z0[b * D + i];
    ostats st;
    tsit5_solve(&m, mode, u, ys, eps, n, t_from, t_to, &st, NULL);
    for (int64_t b = 0; b < n; ++b)
        for (int i = 0; i < nv; ++i) xs_out[b * nv + i] = u[b * R + i];
    copy_stats(stats, &st);
    free(u);
    return st.retcode;
}

/* ------------------------------------------------------------------------------------
 * Discrete adjoint ("discretise-then-optimise") of the TrainMode solve: exact reverse mode
 * of the accepted Tsit5 steps with the accepted dt_n frozen.  The reference
 * obtains its gradient from Enzyme + SciMLSensitivity (call site
 * src/exts/mlj_ext/core_icnf.jl:59-63), which is unpinned; this is the north_star's choice.
 * ------------------------------------------------------------------------------------ */
typedef struct {
    float* gb[CNF_MAX_LAYERS + 1]; /* tangent-like sweep of the eps-chain adjoint */
    float* db[CNF_MAX_LAYERS + 1]; /* adjoint of d_l = 1 - h_l^2 */
    float* hb;                     /* adjoint of h_l (current layer) */
    float* hb2;
    float* pb;
    float* qb;
    float* store;
} obscratch;

static void obscratch_alloc(obscratch* s, const omodel* m) {
    size_t per = (size_t)m->maxd * TS;
    size_t n = (size_t)(2 * (CNF_MAX_LAYERS + 1) + 4) * per;
    s->store = (float*)aligned_alloc(64, n * sizeof(float));
    float* p = s->store;
    for (int l = 0; l <= CNF_MAX_LAYERS; ++l) { s->gb[l] = p; p += per; }
    for (int l = 0; l <= CNF_MAX_LAYERS; ++l) { s->db[l] = p; p += per; }
    s->hb = p; p += per;
    s->hb2 = p; p += per;
    s->pb = p; p += per;
    s->qb = p;
}

/* gW[o,i] += sum_s a[o][s] * b[i][s]   for i < ncols */
static void outer_acc(double* gW, int dout, int ncols, const float* a, const float* b) {
    for (int i = 0; i < ncols; ++i) {
        const float* bi = b + i * TS;
        for (int o = 0; o < dout; ++o) {
            const float* ao = a + o * TS;
            float acc = 0.0f;
            for (int s = 0; s < TS; ++s) acc = fmaf(ao[s], bi[s], acc);
            gW[(size_t)i * dout + o] += (double)acc;
        }
    }
}

/* VJP of the TrainMode augmented_f at stage input s (rows 0:D of u), cotangent
 * (a[D x B], alpha, beta, gamma) on (zdot, ldot, Edot, ndot).  Writes zbar (D x B, stride D)
 * and accumulates parameter gradients into gth (layout of theta). */
/* unit >= 0 (TestMode, exact trace): eps is the unit vector e_unit — the trace row is
 * -sum_i e_i^T J e_i (src/icnf.jl:148-164), so its VJP is the sum over i of the Hutchinson VJP with
 * eps = e_i; the zdot cotangent abar is applied with unit 0 only and zbar accumulates over units. */
static void rhs_vjp_tile(const omodel* m, oscratch* sc, obscratch* bs, const float* u, int R, int unit,
                         const float* ys, const float* eps, const float* abar, float alpha,
                         float beta, float gamma, float* zbar, double* gth, int64_t b0, int ns) {
    const int D = m->D, L = m->L;
    const int jvp = m->cfg.trace_mode == CNF_TRACE_JVP;
    load_h0(m, sc, u, R, ys, b0, ns);
    nn_forward(m, sc);
    float* e = sc->t0;
    for (int i = 0; i < D; ++i)
        for (int s = 0; s < TS; ++s)
            e[i * TS + s] = (s < ns) ? (unit >= 0 ? (i == unit ? 1.0f : 0.0f) : eps[(b0 + s) * D + i]) : 0.0f;
    if (jvp) jvp_chain(m, sc, e); else vjp_chain(m, sc, e);
    const float* y = sc->h[L];
    const float* v = jvp ? sc->g[L] : sc->q[1];
    float ny[TS], nv[TS];
    for (int s = 0; s < TS; ++s) ny[s] = 0.0f, nv[s] = 0.0f;
    for (int i = 0; i < D; ++i)
        for (int s = 0; s < TS; ++s) {
            ny[s] = fmaf(y[i * TS + s], y[i * TS + s], ny[s]);
            nv[s] = fmaf(v[i * TS + s], v[i * TS + s], nv[s]);
        }
    float* ybar = sc->t1;
    float* vbar = sc->t2;
    for (int s = 0; s < TS; ++s) {
        int live = s < ns;
        float nys = sqrtf(ny[s]), nvs = sqrtf(nv[s]);
        float cb = (live && (m->cfg.flags & CNF_FLAG_NORM_Z) && nys > 0.0f) ? beta / nys : 0.0f;
        float cg = (live && (m->cfg.flags & CNF_FLAG_NORM_J) && nvs > 0.0f) ? gamma / nvs : 0.0f;
        float ca = live ? alpha : 0.0f;
        for (int i = 0; i < D; ++i) {
            float ab = (live && unit <= 0) ? abar[(b0 + s) * D + i] : 0.0f;
            ybar[i * TS + s] = fmaf(cb, y[i * TS + s], ab);
            vbar[i * TS + s] = fmaf(cg, v[i * TS + s], -ca * e[i * TS + s]);
        }
    }
    const float* th = m->theta;
    if (!jvp) {
        /* adjoint of the reverse chain runs forward: gb_l = W_l qb_l */
        const float* qb = vbar; /* qb_1 = [vbar; 0] */
        for (int l = 1; l <= L; ++l) {
            int din = m->cfg.dims[l - 1], dout = m->cfg.dims[l];
            int used = (l == 1) ? D : din;
            dense_mul(th + m->w_off[l - 1], NULL, dout, used, qb, bs->gb[l]);
            outer_acc(gth + m->w_off[l - 1], dout, used, sc->g[l], qb);
            const float* hl = H_DER(m, sc, l);
            if (l < L) {
                for (int k = 0; k < dout * TS; ++k) {
                    bs->db[l][k] = bs->gb[l][k] * sc->q[l + 1][k];
                    bs->qb[k] = bs->gb[l][k] * fmaf(-hl[k], hl[k], 1.0f);
                }
                /* next qb must not alias this layer's output: ping-pong through hb2 */
                memcpy(bs->hb2, bs->qb, sizeof(float) * (size_t)dout * TS);
                qb = bs->hb2;
            } else {
                for (int k = 0; k < dout * TS; ++k) bs->db[l][k] = bs->gb[l][k] * e[k];
            }
        }
    } else {
        /* adjoint of the tangent chain runs backward, merged below; precompute db_l, and the
         * W-gradient term rb_l (x) tau_{l-1} */
        float* taub = bs->hb2;
        for (int k = 0; k < D * TS; ++k) taub[k] = vbar[k];
        for (int l = L; l >= 1; --l) {
            int din = m->cfg.dims[l - 1], dout = m->cfg.dims[l];
            int used = (l == 1) ? D : din;
            const float* hl = H_DER(m, sc, l);
            float* rb = bs->gb[l];
            for (int k = 0; k < dout * TS; ++k) {
                rb[k] = taub[k] * fmaf(-hl[k], hl[k], 1.0f);
                bs->db[l][k] = taub[k] * sc->q[l][k];
            }
            outer_acc(gth + m->w_off[l - 1], dout, used, rb, (l == 1) ? e : sc->g[l - 1]);
            if (l > 1) dense_mulT(th + m->w_off[l - 1], dout, din, rb, taub);
        }
    }
    /* standard backward through the forward chain with the extra -2 h db term */
    float* hb = bs->hb;
    for (int k = 0; k < D * TS; ++k) hb[k] = ybar[k];
    for (int l = L; l >= 1; --l) {
        int din = m->cfg.dims[l - 1], dout = m->cfg.dims[l];
        const float* hl = H_DER(m, sc, l);
        for (int k = 0; k < dout * TS; ++k) {
            float hbk = fmaf(-2.0f * hl[k], bs->db[l][k], hb[k]);
            bs->pb[k] = hbk * fmaf(-hl[k], hl[k], 1.0f);
        }
        outer_acc(gth + m->w_off[l - 1], dout, din, bs->pb, sc->h[l - 1]);
        if (!((m->cfg.flags & CNF_FLAG_LAST_LINEAR) && l == L))
            for (int o = 0; o < dout; ++o) {
                float acc = 0.0f;
                for (int s = 0; s < TS; ++s) acc += bs->pb[o * TS + s];
                gth[m->b_off[l - 1] + o] += (double)acc;
            }
        int used = (l == 1) ? D : din;
        dense_mulT(th + m->w_off[l - 1], dout, used, bs->pb, hb);
    }
    for (int s = 0; s < ns; ++s)
        for (int i = 0; i < D; ++i) {
            if (unit > 0) zbar[(b0 + s) * D + i] += hb[i * TS + s];
            else zbar[(b0 + s) * D + i] = hb[i * TS + s];
        }
}

static void rhs_vjp_all(const omodel* m, const float* u, int R, int unit, const float* ys, const float* eps,
                        const float* abar, float alpha, float beta, float gamma, float* zbar,
                        double* gth, int64_t nparams, int64_t B) {
    int64_t ntile = (B + TS - 1) / TS;
#pragma omp parallel
    {
        oscratch sc;
        obscratch bs;
        oscratch_alloc(&sc, m);
        obscratch_alloc(&bs, m);
        double* gl = (double*)calloc((size_t)nparams, sizeof(double));
#pragma omp for schedule(static)
        for (int64_t t = 0; t < ntile; ++t) {
            int64_t b0 = t * TS;
            int ns = (int)((B - b0) < TS ? (B - b0) : TS);
            rhs_vjp_tile(m, &sc, &bs, u, R, unit, ys, eps, abar, alpha, beta, gamma, zbar, gl, b0, ns);
        }
#pragma omp critical
        for (int64_t p = 0; p < nparams; ++p) gth[p] += gl[p];
        free(gl);
        free(bs.store);
        oscratch_free(&sc);
    }
}

/* loss(icnf, TrainMode(), xs[, ys], ps, st) = mean(-logp + l1 E + l2 n + l3 A)
 * (src/icnf.jl:481-502), loss(icnf, TestMode(), ...) = -mean(logp) (src/base_icnf.jl:489-508),
 * and the discrete gradient w.r.t. theta and xs. */
int32_t cnf_oracle_loss_grad_mode(const cnf_config* cfg, const float* theta, int64_t nparams,
                                  const int64_t* w_off, const int64_t* b_off, int32_t mode, const float* xs,
                                  const float* ys, const float* eps, int64_t B, float t0, float t1,
                                  float* loss, float* grad_theta, float* grad_x, cnf_stats* stats) {
    omodel m;
    omodel_init(&m, cfg, theta, w_off, b_off);
    const int train = mode == CNF_MODE_TRAIN;
    const int D = m.D, R = D + (train ? 3 : 1), nv = cfg->nvars;
    const int64_t nel = (int64_t)R * B;
    float* u = assemble_u0(&m, R, xs, B);
    ostats st;
    otape tape;
    memset(&tape, 0, sizeof(tape));
    int want_grad = grad_theta || grad_x;
    tsit5_solve(&m, mode, u, ys, eps, B, t0, t1, &st, want_grad ? &tape : NULL);
    float* logp = (float*)malloc(sizeof(float) * (size_t)B);
    float* regs = (float*)malloc(sizeof(float) * 3 * (size_t)B);
    final_split(&m, mode, u, B, logp, regs, NULL);
    double acc = 0.0;
    for (int64_t b = 0; b < B; ++b)
        acc += train ? (double)(-logp[b] + cfg->lambda1 * regs[b * 3] + cfg->lambda2 * regs[b * 3 + 1] +
                                cfg->lambda3 * regs[b * 3 + 2])
                     : (double)(-logp[b]);
    *loss = (float)(acc / (double)B);
    copy_stats(stats, &st);
    if (want_grad) {
        const float invB = 1.0f / (float)B;
        const int useA = train && (cfg->flags & CNF_FLAG_NORM_Z_AUG) && (cfg->flags & CNF_FLAG_AUGMENTED);
        float* ubar = (float*)malloc(sizeof(float) * (size_t)D * (size_t)B);
        for (int64_t b = 0; b < B; ++b) {
            const float* ub = u + b * R;
            float za = 0.0f;
            for (int i = nv; i < D; ++i) za = fmaf(ub[i], ub[i], za);
            float cA = (useA && za > 0.0f) ? cfg->lambda3 / sqrtf(za) : 0.0f;
            for (int i = 0; i < D; ++i)
                ubar[b * D + i] = (ub[i] + (i >= nv ? cA * ub[i] : 0.0f)) * invB;
        }
        const float c_l = invB, c_E = train ? cfg->lambda1 * invB : 0.0f, c_n = train ? cfg->lambda2 * invB : 0.0f;
        double* gth = (double*)calloc((size_t)nparams, sizeof(double));
        float* k[6];
        float* s[6];
        float* kb[6];
        for (int j = 0; j < 6; ++j) {
            k[j] = (float*)malloc(sizeof(float) * (size_t)nel);
            s[j] = (float*)malloc(sizeof(float) * (size_t)nel);
            kb[j] = (float*)malloc(sizeof(float) * (size_t)D * (size_t)B);
        }
        float* sbar = (float*)malloc(sizeof(float) * (size_t)D * (size_t)B);
        const int64_t nz = (int64_t)D * B;
        for (int64_t n = tape.n - 1; n >= 0; --n) {
            const float dt = tape.dt[n];
            memcpy(s[0], tape.u[n], sizeof(float) * (size_t)nel);
            rhs_all(&m, mode, s[0], ys, eps, k[0], B);
            for (int i = 1; i < 6; ++i) { /* stage checkpointing: recompute the stage inputs */
                rk_combine(s[i], s[0], k, A_[i], i, dt, nel);
                rhs_all(&m, mode, s[i], ys, eps, k[i], B);
            }
            for (int j = 0; j < 6; ++j)
                for (int64_t e = 0; e < nz; ++e) kb[j][e] = dt * A_[6][j] * ubar[e];
            for (int i = 5; i >= 0; --i) {
                float w = dt * A_[6][i];
                if (train) {
                    rhs_vjp_all(&m, s[i], R, -1, ys, eps, kb[i], w * c_l, w * c_E, w * c_n, sbar, gth, nparams, B);
                } else {
                    for (int unit = 0; unit < D; ++unit)
                        rhs_vjp_all(&m, s[i], R, unit, ys, eps, kb[i], w * c_l, 0.0f, 0.0f, sbar, gth, nparams, B);
                }
                for (int64_t e = 0; e < nz; ++e) ubar[e] += sbar[e];
                for (int j = 0; j < i; ++j) {
                    float c = dt * A_[i][j];
                    for (int64_t e = 0; e < nz; ++e) kb[j][e] = fmaf(c, sbar[e], kb[j][e]);
                }
            }
        }
        if (grad_theta)
            for (int64_t p = 0; p < nparams; ++p) grad_theta[p] = (float)gth[p];
        if (grad_x)
            for (int64_t b = 0; b < B; ++b)
                for (int i = 0; i < nv; ++i) grad_x[b * nv + i] = ubar[b * D + i];
        for (int j = 0; j < 6; ++j) { free(k[j]); free(s[j]); free(kb[j]); }
        free(sbar); free(gth); free(ubar);
    }
    tape_free(&tape);
    free(logp); free(regs); free(u);
    return st.retcode;
}

int32_t cnf_oracle_loss_grad(const cnf_config* cfg, const float* theta, int64_t nparams,
                             const int64_t* w_off, const int64_t* b_off, const float* xs,
                             const float* ys, const float* eps, int64_t B, float t0, float t1,
                             float* loss, float* grad_theta, float* grad_x, cnf_stats* stats) {
    return cnf_oracle_loss_grad_mode(cfg, theta, nparams, w_off, b_off, CNF_MODE_TRAIN, xs, ys, eps, B, t0, t1,
                                     loss, grad_theta, grad_x, stats);
}

int32_t cnf_oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* torchrun exports OMP_NUM_THREADS=1 to every rank; the timed CPU legs of bench.py ask for all host cores explicitly */
void cnf_oracle_set_num_threads(int32_t n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* ---- "next" row N1: the generator behind cnfb200_seed_rng, restated on the CPU.
 * Philox4x32-10 (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as 1, 2, 3", SC'11;
 * the Random123 known-answer vectors pin it in tests/test_oracle.py), counter =
 * (lo32(sample), hi32(sample), row group, draw), key = (lo32(seed), hi32(seed)); 23-bit uniforms
 * u = ((x >> 9) + 1/2) 2^-23 and Box-Muller pairs (x0, x1) -> rows 4g, 4g+1; (x2, x3) -> 4g+2, 4g+3.
 * Replaces randn!(rng, eps) / rand(rng, basedist, n) of src/base_icnf.jl:277-278, 365-372 when the
 * caller opts in; it is NOT Julia's stream. */
void cnf_oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static float philox_u01(uint32_t x) { return ((float)(x >> 9) + 0.5f) * 1.1920928955078125e-07f; }

/* out: B x rows, sample-major (the caller's D x B column-major layout) */
void cnf_oracle_philox_normal(uint64_t seed, uint64_t sample_offset, uint32_t draw, int32_t rows, int64_t B, float* out) {
    const uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    const double two_pi = 6.283185307179586476925286766559;
    for (int64_t b = 0; b < B; ++b) {
        const uint64_t gs = (uint64_t)b + sample_offset;
        for (int g = 0; 4 * g < rows; ++g) {
            const uint32_t ctr[4] = {(uint32_t)gs, (uint32_t)(gs >> 32), (uint32_t)g, draw};
            uint32_t x[4];
            cnf_oracle_philox4x32_10(ctr, key, x);
            float n[4];
            for (int h = 0; h < 2; ++h) {
                const float r = sqrtf(-2.0f * logf(philox_u01(x[2 * h])));
                const double th = two_pi * (double)philox_u01(x[2 * h + 1]);
                n[2 * h] = r * (float)cos(th);
                n[2 * h + 1] = r * (float)sin(th);
            }
            for (int j = 0; j < 4; ++j)
                if (4 * g + j < rows) out[b * rows + 4 * g + j] = n[j];
        }
    }
}
