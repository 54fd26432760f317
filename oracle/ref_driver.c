/* oracle/ref_driver.c -- closed loop of the reference on the UNMODIFIED compiled acados core
 * (TEST INFRASTRUCTURE; links oracle/_ref/libacados_ref.a built by oracle/ref.mk from /root/reference).
 *
 * The reference builds its solver by Python codegen (casadi + tera templates + make), none of which
 * can run here.  This driver does by hand what the generated `acados_solver_<name>.c` does for the
 * one OCP the reference formulates, call for call:
 *   create sequence   T/c_templates_tera/acados_solver.in.c:2795-2849 (plan :176-220, dims :223-406,
 *                     opts :2286-2738, initial guess :2744-2777, nlp_in :949-2280)
 *   option values     R/trajectory_optimizer.py:60-70 on top of T/acados_ocp_options.py:45-165 and
 *                     T/acados_ocp.py:1122-1264
 *   closed loop       R/simulator.py:203-226 (set lbx/ubx at stage 0, solve, get u0, stats, plant RK4)
 * The CasADi-generated model functions are replaced by the hand restatement in ur10_model.c, wrapped
 * as acados `external_function_generic` objects (AC/utils/external_function_generic.h:86-94).
 * Everything between those two ends -- SQP / SQP_RTI, merit line search, NLS cost, BGH bounds,
 * x0 elimination, HPIPM, BLASFEO -- is the reference's own compiled code.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "acados/utils/external_function_generic.h"
#include "acados/utils/math.h"
#include "acados/utils/print.h"
#include "acados_c/external_function_interface.h"
#include "acados_c/ocp_nlp_interface.h"
#include "blasfeo_d_aux.h"
#include "blasfeo_d_aux_ext_dep.h"

#include "acados/ocp_nlp/ocp_nlp_common.h"
#include "acados/ocp_qp/ocp_qp_hpipm.h"
#include "acados/ocp_qp/ocp_qp_xcond_solver.h"
#include "hpipm_d_ocp_qp_ipm.h"

#include "rmpc_problem.h"

#define NX 12
#define NU 6
#define NY 18

/* ------------------------------------------------------------------ external function wrappers */
typedef struct
{
    external_function_generic base; /* must be first: acados calls base.evaluate(self, ...) */
    const ur10_params *P;
    int kind;
    const double *term_P; /* K_TERM_*: 12 x 12 symmetric, row-major */
} ur10_ext_fun;

enum { K_DYN_FUN, K_DYN_FUN_JAC, K_COST_FUN, K_COST_FUN_JAC, K_TERM_FUN, K_TERM_FUN_JAC };

static void read_vec(ext_fun_arg_t t, void *a, int n, double *out)
{
    if (t == BLASFEO_DVEC_ARGS)
    {
        struct blasfeo_dvec_args *v = a;
        for (int i = 0; i < n; i++) out[i] = BLASFEO_DVECEL(v->x, v->xi + i);
    }
    else if (t == BLASFEO_DVEC)
    {
        struct blasfeo_dvec *v = a;
        for (int i = 0; i < n; i++) out[i] = BLASFEO_DVECEL(v, i);
    }
    else if (t == COLMAJ)
    {
        memcpy(out, a, n * sizeof(double));
    }
    else
    {
        fprintf(stderr, "ref_driver: unsupported input arg type %d\n", (int) t);
        exit(1);
    }
}

static void write_vec(ext_fun_arg_t t, void *a, int n, const double *in)
{
    if (t == BLASFEO_DVEC_ARGS)
    {
        struct blasfeo_dvec_args *v = a;
        for (int i = 0; i < n; i++) BLASFEO_DVECEL(v->x, v->xi + i) = in[i];
    }
    else if (t == BLASFEO_DVEC)
    {
        struct blasfeo_dvec *v = a;
        for (int i = 0; i < n; i++) BLASFEO_DVECEL(v, i) = in[i];
    }
    else if (t == COLMAJ)
    {
        memcpy(a, in, n * sizeof(double));
    }
    else
    {
        fprintf(stderr, "ref_driver: unsupported output vec type %d\n", (int) t);
        exit(1);
    }
}

static void write_mat(ext_fun_arg_t t, void *a, int m, int n, const double *in /* colmaj ld=m */)
{
    if (t == BLASFEO_DMAT_ARGS)
    {
        struct blasfeo_dmat_args *M = a;
        for (int j = 0; j < n; j++)
            for (int i = 0; i < m; i++) BLASFEO_DMATEL(M->A, M->ai + i, M->aj + j) = in[i + m * j];
    }
    else if (t == BLASFEO_DMAT)
    {
        struct blasfeo_dmat *M = a;
        for (int j = 0; j < n; j++)
            for (int i = 0; i < m; i++) BLASFEO_DMATEL(M, i, j) = in[i + m * j];
    }
    else if (t == IGNORE_ARGUMENT)
    {
    }
    else
    {
        fprintf(stderr, "ref_driver: unsupported output mat type %d\n", (int) t);
        exit(1);
    }
}

static void ur10_ext_evaluate(void *self_, ext_fun_arg_t *tin, void **in, ext_fun_arg_t *tout, void **out)
{
    ur10_ext_fun *self = self_;
    double x[NX], u[NU];
    read_vec(tin[0], in[0], NX, x);
    if (self->kind == K_TERM_FUN || self->kind == K_TERM_FUN_JAC)
    { /* h_e(x) = x' P x and its gradient 2 P x at stage N (nu = 0): the *_constr_h_e_fun[_jac_uxt_zt] pair of
       * T/casadi_function_generation.py for con_h_expr_e = x' P x */
        double h = 0.0, g[NX];
        for (int r = 0; r < NX; r++)
        {
            double a = 0.0;
            for (int k = 0; k < NX; k++) a += self->term_P[NX * r + k] * x[k];
            g[r] = 2.0 * a;
            h += x[r] * a;
        }
        write_vec(tout[0], out[0], 1, &h);
        if (self->kind == K_TERM_FUN_JAC) write_mat(tout[1], out[1], NX, 1, g);
        return;
    }
    read_vec(tin[1], in[1], NU, u);
    switch (self->kind)
    {
        case K_DYN_FUN:
        case K_DYN_FUN_JAC:
        {
            double xn[NX];
            ur10_dyn(self->P, x, u, xn);
            write_vec(tout[0], out[0], NX, xn);
            if (self->kind == K_DYN_FUN_JAC)
            {
                double BAbt[18 * 12];
                ur10_dyn_jac_BAbt(self->P, BAbt);
                write_mat(tout[1], out[1], NU + NX, NX, BAbt);
            }
            break;
        }
        case K_COST_FUN:
        {
            double y[NY];
            ur10_cost_y(self->P, x, u, y);
            write_vec(tout[0], out[0], NY, y);
            break;
        }
        case K_COST_FUN_JAC:
        {
            double y[NY], Jt[18 * 18];
            ur10_cost_y_jac(self->P, x, u, y, Jt);
            write_vec(tout[0], out[0], NY, y);
            write_mat(tout[1], out[1], NU + NX, NY, Jt);
            /* out[2] = dy/dz (ny x nz, nz = 0): nothing to write */
            break;
        }
    }
}

static size_t ur10_ext_ws_req(void *self) { (void) self; return 0; }
static void ur10_ext_ws_set(void *self, void *w) { (void) self; (void) w; }

static void ur10_ext_init(ur10_ext_fun *f, const ur10_params *P, int kind)
{
    memset(f, 0, sizeof(*f));
    f->base.evaluate = &ur10_ext_evaluate;
    f->base.get_external_workspace_requirement = &ur10_ext_ws_req;
    f->base.set_external_workspace = &ur10_ext_ws_set;
    f->P = P;
    f->kind = kind;
}

/* ------------------------------------------------------------------ capsule */
typedef struct
{
    rmpc_problem prob;
    ocp_nlp_plan_t *plan;
    ocp_nlp_config *config;
    ocp_nlp_dims *dims;
    void *opts;
    ocp_nlp_in *in;
    ocp_nlp_out *out;
    ocp_nlp_solver *solver;
    ur10_ext_fun dyn_fun, dyn_fun_jac, cost_fun, cost_fun_jac, term_fun, term_fun_jac;
} rmpc_ref;

void rmpc_ref_destroy(rmpc_ref *c)
{
    if (!c) return;
    /* same order as <name>_acados_free, acados_solver.in.c */
    if (c->opts) ocp_nlp_solver_opts_destroy(c->opts);
    if (c->in) ocp_nlp_in_destroy(c->in);
    if (c->out) ocp_nlp_out_destroy(c->out);
    if (c->solver) ocp_nlp_solver_destroy(c->solver);
    if (c->dims) ocp_nlp_dims_destroy(c->dims);
    if (c->config) ocp_nlp_config_destroy(c->config);
    if (c->plan) ocp_nlp_plan_destroy(c->plan);
    free(c);
}

rmpc_ref *rmpc_ref_create(const rmpc_problem *prob, const double x0[12])
{
    rmpc_ref *c = calloc(1, sizeof(rmpc_ref));
    c->prob = *prob;
    const rmpc_problem *p = &c->prob;
    const int N = p->N;

    /* 1) plan (acados_solver.in.c:176-220) */
    c->plan = ocp_nlp_plan_create(N);
    c->plan->nlp_solver = p->solver_type == RMPC_SQP_RTI ? SQP_RTI : SQP;
    c->plan->ocp_qp_solver_plan.qp_solver = PARTIAL_CONDENSING_HPIPM;
    c->plan->relaxed_ocp_qp_solver_plan.qp_solver = PARTIAL_CONDENSING_HPIPM;
    for (int i = 0; i < N; i++) c->plan->nlp_cost[i] = NONLINEAR_LS; /* cost_type_0 copies cost_type, T/acados_ocp.py:200-201 */
    c->plan->nlp_cost[N] = LINEAR_LS;                                /* default cost_type_e with ny_e = 0 */
    for (int i = 0; i < N; i++)
    {
        c->plan->nlp_dynamics[i] = DISCRETE_MODEL;
        c->plan->sim_solver_plan[i].sim_solver = INVALID_SIM_SOLVER;
    }
    for (int i = 0; i <= N; i++) c->plan->nlp_constraints[i] = BGH;
    c->plan->regularization = NO_REGULARIZE;
    c->plan->globalization = MERIT_BACKTRACKING;
    c->config = ocp_nlp_config_create(*c->plan);

    /* 2) dims (acados_solver.in.c:223-406) */
    int *mem = calloc((size_t) 8 * (N + 1), sizeof(int));
    int *nx = mem, *nu = mem + (N + 1), *nz = mem + 2 * (N + 1), *ns = mem + 3 * (N + 1),
        *np = mem + 4 * (N + 1);
    for (int i = 0; i <= N; i++) { nx[i] = NX; nu[i] = NU; }
    nu[N] = 0;
    c->dims = ocp_nlp_dims_create(c->config);
    ocp_nlp_dims_set_opt_vars(c->config, c->dims, "nx", nx);
    ocp_nlp_dims_set_opt_vars(c->config, c->dims, "nu", nu);
    ocp_nlp_dims_set_opt_vars(c->config, c->dims, "nz", nz);
    ocp_nlp_dims_set_opt_vars(c->config, c->dims, "ns", ns);
    ocp_nlp_dims_set_opt_vars(c->config, c->dims, "np", np);
    ocp_nlp_dims_set_global(c->config, c->dims, "np_global", 0);
    ocp_nlp_dims_set_global(c->config, c->dims, "n_global_data", 0);
    for (int i = 0; i <= N; i++)
    {
        int nbx = (i == 0) ? NX : (i < N ? 6 : 0);
        int nbu = (i < N) ? NU : 0;
        int zero = 0;
        int nbxe = (i == 0) ? NX : 0;
        ocp_nlp_dims_set_constraints(c->config, c->dims, i, "nbx", &nbx);
        ocp_nlp_dims_set_constraints(c->config, c->dims, i, "nbu", &nbu);
        ocp_nlp_dims_set_constraints(c->config, c->dims, i, "nsbx", &zero);
        ocp_nlp_dims_set_constraints(c->config, c->dims, i, "nsbu", &zero);
        ocp_nlp_dims_set_constraints(c->config, c->dims, i, "ng", &zero);
        ocp_nlp_dims_set_constraints(c->config, c->dims, i, "nsg", &zero);
        ocp_nlp_dims_set_constraints(c->config, c->dims, i, "nbxe", &nbxe);
    }
    {
        int ny = NY, zero = 0;
        for (int i = 0; i < N; i++) ocp_nlp_dims_set_cost(c->config, c->dims, i, "ny", &ny);
        for (int i = 0; i <= N; i++)
        {
            int nh = (i == N && p->term) ? 1 : 0; /* nh_e = 1: terminal ellipsoid */
            ocp_nlp_dims_set_constraints(c->config, c->dims, i, "nh", &nh);
            ocp_nlp_dims_set_constraints(c->config, c->dims, i, "nsh", &zero);
        }
        ocp_nlp_dims_set_cost(c->config, c->dims, N, "ny", &zero);
    }
    free(mem);

    /* 3) opts (acados_solver.in.c:2286-2738), in the template's order */
    c->opts = ocp_nlp_solver_opts_create(c->config, c->dims);
    void *o = c->opts;
    ocp_nlp_config *cf = c->config;
    {
        int fixed_hess = 0;
        ocp_nlp_solver_opts_set(cf, o, "fixed_hess", &fixed_hess);
        double alpha_min = 0.05, alpha_red = 0.7; /* T/acados_ocp.py:1222-1227, T/acados_ocp_options.py:128 */
        ocp_nlp_solver_opts_set(cf, o, "globalization_alpha_min", &alpha_min);
        ocp_nlp_solver_opts_set(cf, o, "globalization_alpha_reduction", &alpha_red);
        int suff = 0, soc = 0;
        double eps_suff = 1e-4;
        ocp_nlp_solver_opts_set(cf, o, "globalization_line_search_use_sufficient_descent", &suff);
        ocp_nlp_solver_opts_set(cf, o, "globalization_use_SOC", &soc);
        ocp_nlp_solver_opts_set(cf, o, "globalization_eps_sufficient_descent", &eps_suff);
        int sens = 0;
        ocp_nlp_solver_opts_set(cf, o, "with_solution_sens_wrt_params", &sens);
        ocp_nlp_solver_opts_set(cf, o, "with_value_sens_wrt_params", &sens);
        double t_lam_min = 1e-9;
        ocp_nlp_solver_opts_set(cf, o, "solution_sens_qp_t_lam_min", &t_lam_min);
        int full_step_dual = 0;
        ocp_nlp_solver_opts_set(cf, o, "globalization_full_step_dual", &full_step_dual);
        int ws_first = 1; /* nlp_solver_warm_start_first_qp = True, R/trajectory_optimizer.py:65 */
        ocp_nlp_solver_opts_set(cf, o, "warm_start_first_qp", &ws_first);
        /* warm_start_first_qp_from_nlp is False in python => template does not set it */
        double lm = 0.0;
        ocp_nlp_solver_opts_set(cf, o, "levenberg_marquardt", &lm);
        int cond_N = (p->cond_N > 0 && p->cond_N < N) ? p->cond_N : N; /* qp_solver_cond_N = None -> N: no condensing */
        ocp_nlp_solver_opts_set(cf, o, "qp_cond_N", &cond_N);
        int ext_qp_res = 0;
        ocp_nlp_solver_opts_set(cf, o, "ext_qp_res", &ext_qp_res);
        bool store_it = false;
        ocp_nlp_solver_opts_set(cf, o, "store_iterates", &store_it);
        if (p->solver_type == RMPC_SQP)
        {
            int lg = 0;
            ocp_nlp_solver_opts_set(cf, o, "log_primal_step_norm", &lg);
            ocp_nlp_solver_opts_set(cf, o, "log_dual_step_norm", &lg);
            double tol_min_step = 0.0;
            ocp_nlp_solver_opts_set(cf, o, "tol_min_step_norm", &tol_min_step);
        }
        ocp_nlp_solver_opts_set(cf, o, "qp_hpipm_mode", "BALANCE");
        /* qp_solver_mu0 = 0 -> not set; */
        int t0_init = 2;
        ocp_nlp_solver_opts_set(cf, o, "qp_t0_init", &t0_init);
        /* tau_min = 0 -> not set */
        if (p->solver_type == RMPC_SQP)
        {
            double t;
            t = p->tol[0]; ocp_nlp_solver_opts_set(cf, o, "tol_stat", &t);
            t = p->tol[1]; ocp_nlp_solver_opts_set(cf, o, "tol_eq", &t);
            t = p->tol[2]; ocp_nlp_solver_opts_set(cf, o, "tol_ineq", &t);
            t = p->tol[3]; ocp_nlp_solver_opts_set(cf, o, "tol_comp", &t);
            int max_iter = p->max_iter;
            ocp_nlp_solver_opts_set(cf, o, "max_iter", &max_iter);
            bool alm = false;
            ocp_nlp_solver_opts_set(cf, o, "with_adaptive_levenberg_marquardt", &alm);
            double alm_lam = 5.0, alm_mu_min = 1e-16, alm_mu0 = 1e-3, alm_obj = 2.0;
            ocp_nlp_solver_opts_set(cf, o, "adaptive_levenberg_marquardt_lam", &alm_lam);
            ocp_nlp_solver_opts_set(cf, o, "adaptive_levenberg_marquardt_mu_min", &alm_mu_min);
            ocp_nlp_solver_opts_set(cf, o, "adaptive_levenberg_marquardt_mu0", &alm_mu0);
            ocp_nlp_solver_opts_set(cf, o, "adaptive_levenberg_marquardt_obj_scalar", &alm_obj);
            bool eval_res_max = false;
            ocp_nlp_solver_opts_set(cf, o, "eval_residual_at_max_iter", &eval_res_max);
            double ub_eig = 1e5, lb_grad = 1e-4;
            ocp_nlp_solver_opts_set(cf, o, "qpscaling_ub_max_abs_eig", &ub_eig);
            ocp_nlp_solver_opts_set(cf, o, "qpscaling_lb_norm_inf_grad_obj", &lb_grad);
            qpscaling_scale_objective_type so = NO_OBJECTIVE_SCALING;
            ocp_nlp_solver_opts_set(cf, o, "qpscaling_scale_objective", &so);
            ocp_nlp_qpscaling_constraint_type sc = NO_CONSTRAINT_SCALING;
            ocp_nlp_solver_opts_set(cf, o, "qpscaling_scale_constraints", &sc);
            ocp_nlp_qp_tol_strategy_t strat = FIXED_QP_TOL;
            ocp_nlp_solver_opts_set(cf, o, "nlp_qp_tol_strategy", &strat);
            double red = 1e-1, safety = 0.1, mstat = 1e-9, meq = 1e-10, mineq = 1e-10, mcomp = 1e-11;
            ocp_nlp_solver_opts_set(cf, o, "nlp_qp_tol_reduction_factor", &red);
            ocp_nlp_solver_opts_set(cf, o, "nlp_qp_tol_safety_factor", &safety);
            ocp_nlp_solver_opts_set(cf, o, "nlp_qp_tol_min_stat", &mstat);
            ocp_nlp_solver_opts_set(cf, o, "nlp_qp_tol_min_eq", &meq);
            ocp_nlp_solver_opts_set(cf, o, "nlp_qp_tol_min_ineq", &mineq);
            ocp_nlp_solver_opts_set(cf, o, "nlp_qp_tol_min_comp", &mcomp);
        }
        else
        {
            int as_rti_iter = p->as_rti_iter > 0 ? p->as_rti_iter : 1, as_rti_level = p->as_rti_level_p1 > 0 ? p->as_rti_level_p1 - 1 : 4, rti_log = 0, rti_log_avail = 0;
            ocp_nlp_solver_opts_set(cf, o, "as_rti_iter", &as_rti_iter);
            ocp_nlp_solver_opts_set(cf, o, "as_rti_level", &as_rti_level);
            ocp_nlp_solver_opts_set(cf, o, "rti_log_residuals", &rti_log);
            ocp_nlp_solver_opts_set(cf, o, "rti_log_only_available_residuals", &rti_log_avail);
        }
        bool anderson = false;
        double anderson_thr = 1e1;
        ocp_nlp_solver_opts_set(cf, o, "with_anderson_acceleration", &anderson);
        ocp_nlp_solver_opts_set(cf, o, "anderson_activation_threshold", &anderson_thr);
        int qp_iter_max = p->qp_iter_max;
        ocp_nlp_solver_opts_set(cf, o, "qp_iter_max", &qp_iter_max);
        double qt = p->qp_tol; /* qp_tol sets all four, T/acados_ocp_options.py:1419-1427 */
        ocp_nlp_solver_opts_set(cf, o, "qp_tol_stat", &qt);
        ocp_nlp_solver_opts_set(cf, o, "qp_tol_eq", &qt);
        ocp_nlp_solver_opts_set(cf, o, "qp_tol_ineq", &qt);
        ocp_nlp_solver_opts_set(cf, o, "qp_tol_comp", &qt);
        int qp_ws = 2; /* R/trajectory_optimizer.py:66 */
        ocp_nlp_solver_opts_set(cf, o, "qp_warm_start", &qp_ws);
        int print_level = 0;
        ocp_nlp_solver_opts_set(cf, o, "print_level", &print_level);
        int cond_ric = 1, ric = 1;
        ocp_nlp_solver_opts_set(cf, o, "qp_cond_ric_alg", &cond_ric);
        ocp_nlp_solver_opts_set(cf, o, "qp_ric_alg", &ric);
    }

    /* 4) nlp_out + initial guess x_k = x0, u_k = 0 (acados_solver.in.c:2744-2777) */
    c->out = ocp_nlp_out_create(c->config, c->dims);
    {
        double xu0[NX + NU];
        memset(xu0, 0, sizeof(xu0));
        memcpy(xu0, x0, NX * sizeof(double));
        for (int i = 0; i < N; i++)
        {
            ocp_nlp_out_set(c->config, c->dims, c->out, c->in, i, "x", xu0);
            ocp_nlp_out_set(c->config, c->dims, c->out, c->in, i, "u", xu0 + NX);
        }
        ocp_nlp_out_set(c->config, c->dims, c->out, c->in, N, "x", xu0);
    }

    /* 5) nlp_in */
    c->in = ocp_nlp_in_create(c->config, c->dims);

    /* 6) functions + nlp_in */
    ur10_ext_init(&c->dyn_fun, &c->prob.model, K_DYN_FUN);
    ur10_ext_init(&c->dyn_fun_jac, &c->prob.model, K_DYN_FUN_JAC);
    ur10_ext_init(&c->cost_fun, &c->prob.model, K_COST_FUN);
    ur10_ext_init(&c->cost_fun_jac, &c->prob.model, K_COST_FUN_JAC);
    {
        double ts = p->dt;
        for (int i = 0; i < N; i++) ocp_nlp_in_set(c->config, c->dims, c->in, i, "Ts", &ts);
        for (int i = 0; i <= N; i++)
        {
            double sc = i < N ? p->dt : 1.0; /* cost_scaling = [time_steps, 1.0], T/acados_ocp.py:1122-1123 */
            ocp_nlp_cost_model_set(c->config, c->dims, c->in, i, "scaling", &sc);
        }
        for (int i = 0; i < N; i++)
        {
            ocp_nlp_dynamics_model_set(c->config, c->dims, c->in, i, "disc_dyn_fun", &c->dyn_fun);
            ocp_nlp_dynamics_model_set(c->config, c->dims, c->in, i, "disc_dyn_fun_jac", &c->dyn_fun_jac);
        }
        double yref[NY], W[NY * NY];
        memset(yref, 0, sizeof(yref));
        memset(W, 0, sizeof(W));
        yref[1] = 1.0; /* g_ref = [0,1,0,px_ref,vy_ref], R/trajectory_optimizer.py:109-115 */
        yref[3] = p->px_ref;
        yref[4] = p->vy_ref;
        for (int i = 0; i < 5; i++) W[i + NY * i] = p->w_task[i];
        for (int i = 0; i < 6; i++) W[(5 + i) + NY * (5 + i)] = 2.0 * p->w_u;
        for (int i = 0; i < 6; i++) W[(11 + i) + NY * (11 + i)] = p->w_qddot;
        W[17 + NY * 17] = 0.0; /* w_manip = 0 */
        for (int i = 0; i < N; i++)
        {
            ocp_nlp_cost_model_set(c->config, c->dims, c->in, i, "yref", yref);
            ocp_nlp_cost_model_set(c->config, c->dims, c->in, i, "W", W);
        }
        for (int i = 0; i < N; i++)
        {
            ocp_nlp_cost_model_set(c->config, c->dims, c->in, i, "nls_y_fun", &c->cost_fun);
            ocp_nlp_cost_model_set(c->config, c->dims, c->in, i, "nls_y_fun_jac", &c->cost_fun_jac);
        }
        /* constraints */
        int idx12[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
        double lbx0[NX], ubx0[NX];
        memcpy(lbx0, x0, sizeof(lbx0));
        memcpy(ubx0, x0, sizeof(ubx0));
        ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, 0, "idxbx", idx12);
        ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, 0, "lbx", lbx0);
        ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, 0, "ubx", ubx0);
        ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, 0, "idxbxe", idx12);
        double lbu[NU], ubu[NU], lbx[6], ubx[6];
        memcpy(lbu, p->u_min, sizeof(lbu));
        memcpy(ubu, p->u_max, sizeof(ubu));
        memcpy(lbx, p->q_min, sizeof(lbx));
        memcpy(ubx, p->q_max, sizeof(ubx));
        for (int i = 0; i < N; i++)
        {
            ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, i, "idxbu", idx12);
            ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, i, "lbu", lbu);
            ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, i, "ubu", ubu);
        }
        for (int i = 1; i < N; i++)
        {
            ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, i, "idxbx", idx12);
            ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, i, "lbx", lbx);
            ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, i, "ubx", ubx);
        }
    }

    if (p->term)
    { /* acados_solver.in.c:2109-2116 */
        ur10_ext_init(&c->term_fun, &c->prob.model, K_TERM_FUN);
        ur10_ext_init(&c->term_fun_jac, &c->prob.model, K_TERM_FUN_JAC);
        c->term_fun.term_P = c->prob.term_P;
        c->term_fun_jac.term_P = c->prob.term_P;
        double lh = p->term_lh, uh = p->term_uh;
        ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, N, "nl_constr_h_fun_jac", &c->term_fun_jac);
        ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, N, "nl_constr_h_fun", &c->term_fun);
        ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, N, "lh", &lh);
        ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, N, "uh", &uh);
    }

    /* 7) solver, 8) precompute */
    c->solver = ocp_nlp_solver_create(c->config, c->dims, c->opts, c->in);
    int status = ocp_nlp_precompute(c->solver, c->in, c->out);
    if (status != ACADOS_SUCCESS)
    {
        fprintf(stderr, "ref_driver: ocp_nlp_precompute failed (%d)\n", status);
        rmpc_ref_destroy(c);
        return NULL;
    }
    return c;
}

/* One MPC step exactly as R/simulator.py:210-221 drives the solver. */
int rmpc_ref_step(rmpc_ref *c, const double x[12], double u0[6], int *sqp_iter, int *qp_iter,
                  double res[4], double *cost, double *time_tot)
{
    double xb[NX];
    memcpy(xb, x, sizeof(xb));
    ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, 0, "lbx", xb);
    ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, 0, "ubx", xb);
    int status = ocp_nlp_solve(c->solver, c->in, c->out);
    ocp_nlp_out_get(c->config, c->dims, c->out, 0, "u", u0);
    if (sqp_iter) ocp_nlp_get(c->solver, "sqp_iter", sqp_iter);
    if (qp_iter)
    {
        /* sum of QP iterations over the SQP iterations of this call (statistics table) */
        int it = 0, stat_n = 0, stat_m = 0;
        double *stat = NULL;
        ocp_nlp_get(c->solver, "sqp_iter", &it);
        ocp_nlp_get(c->solver, "stat_n", &stat_n);
        ocp_nlp_get(c->solver, "stat_m", &stat_m);
        ocp_nlp_get(c->solver, "stat", &stat);
        int tot = 0;
        if (c->prob.solver_type == RMPC_SQP_RTI)
            tot = (int) stat[stat_n * 1 + 1]; /* AC/ocp_nlp/ocp_nlp_sqp_rti.c:607 */
        else
            for (int k = 1; k <= it && k < stat_m; k++) tot += (int) stat[stat_n * k + 5]; /* ocp_nlp_sqp.c:701 */
        *qp_iter = tot;
    }
    if (res)
    {
        ocp_nlp_get(c->solver, "res_stat", &res[0]);
        ocp_nlp_get(c->solver, "res_eq", &res[1]);
        ocp_nlp_get(c->solver, "res_ineq", &res[2]);
        ocp_nlp_get(c->solver, "res_comp", &res[3]);
    }
    if (time_tot) ocp_nlp_get(c->solver, "time_tot", time_tot);
    if (cost)
    {
        ocp_nlp_eval_cost(c->solver, c->in, c->out); /* T/acados_ocp_solver.py:1664-1679 */
        ocp_nlp_get(c->solver, "cost_value", cost);
    }
    return status;
}

/* AcadosOcpSolver.cost_set / constraints_set between two solves (T/acados_ocp_solver.py:1863-2035): the reference's own
 * ocp_nlp_cost_model_set / ocp_nlp_constraints_model_set (acados_c/ocp_nlp_interface.c:516-560,637-760).
 * field: "yref" (18), "W" (18 x 18 column-major), "lbu" / "ubu" (6), "lbx" / "ubx" (12 at stage 0, 6 on q at stages 1..N-1). */
int rmpc_ref_model_set(rmpc_ref *c, int stage, const char *field, const double *value)
{
    if (!strcmp(field, "yref") || !strcmp(field, "W"))
        return ocp_nlp_cost_model_set(c->config, c->dims, c->in, stage, field, (void *) value);
    return ocp_nlp_constraints_model_set(c->config, c->dims, c->in, c->out, stage, field, (void *) value);
}

/* the statistics table of the last solve as AcadosOcpSolver.get_stats('statistics') returns it: out[(stat_n + 1) x rows],
 * row 0 = iteration index (ocp_nlp_get "statistics", AC/ocp_nlp/ocp_nlp_sqp.c:960-975); returns rows, *stat_n_out = stat_n */
int rmpc_ref_statistics(rmpc_ref *c, double *out, int max_doubles, int *stat_n_out)
{
    int it = 0, stat_n = 0, stat_m = 0;
    ocp_nlp_get(c->solver, "nlp_iter", &it);
    ocp_nlp_get(c->solver, "stat_n", &stat_n);
    ocp_nlp_get(c->solver, "stat_m", &stat_m);
    const int rows = stat_m < it + 1 ? stat_m : it + 1;
    if (stat_n_out) *stat_n_out = stat_n;
    if ((stat_n + 1) * rows > max_doubles) return -1;
    ocp_nlp_get(c->solver, "statistics", out);
    return rows;
}

/* acados' own timing breakdown of the last solve: time_tot, time_lin, time_qp, time_qp_solver_call, time_qp_xcond, time_glob */
void rmpc_ref_times(rmpc_ref *c, double t[6])
{
    static const char *names[6] = {"time_tot", "time_lin", "time_qp_sol", "time_qp_solver_call", "time_qp_xcond", "time_glob"};
    for (int k = 0; k < 6; k++) ocp_nlp_get(c->solver, names[k], &t[k]);
}

/* opts 'rti_phase' (0 preparation + feedback, 1 preparation, 2 feedback; AC/ocp_nlp/ocp_nlp_sqp_rti.c:1140-1170), as
 * AcadosOcpSolver.options_set('rti_phase', .) does (T/acados_ocp_solver.py:2326-2340). */
void rmpc_ref_set_rti_phase(rmpc_ref *c, int phase)
{
    ocp_nlp_solver_opts_set(c->config, c->opts, "rti_phase", &phase);
}

/* Preparation phase only: the solve call without touching the stage-0 bounds. */
int rmpc_ref_prepare(rmpc_ref *c)
{
    rmpc_ref_set_rti_phase(c, 1);
    int status = ocp_nlp_solve(c->solver, c->in, c->out);
    rmpc_ref_set_rti_phase(c, 0);
    return status;
}

void rmpc_ref_get_iterate(rmpc_ref *c, double *xtraj, double *utraj, double *pitraj)
{
    const int N = c->prob.N;
    for (int i = 0; i <= N; i++)
        if (xtraj) ocp_nlp_out_get(c->config, c->dims, c->out, i, "x", xtraj + NX * i);
    for (int i = 0; i < N; i++)
    {
        if (utraj) ocp_nlp_out_get(c->config, c->dims, c->out, i, "u", utraj + NU * i);
        if (pitraj) ocp_nlp_out_get(c->config, c->dims, c->out, i, "pi", pitraj + NX * i);
    }
}

/* lam of stage i (2*nb doubles, [lower; upper], bound order [u; x]) */
void rmpc_ref_get_lam(rmpc_ref *c, int stage, double *lam)
{
    ocp_nlp_out_get(c->config, c->dims, c->out, stage, "lam", lam);
}

/* Primal shift of the initial guess through the solver's own get / set (what an acados MPC loop does between two solves:
 * for i < N-1: set(i, 'x', get(i+1, 'x')), set(i, 'u', get(i+1, 'u')); set(N-1, 'x', get(N, 'x'))). */
void rmpc_ref_shift(rmpc_ref *c)
{
    const int N = c->prob.N;
    double v[NX];
    for (int i = 0; i + 1 < N; i++)
    {
        ocp_nlp_out_get(c->config, c->dims, c->out, i + 1, "x", v);
        ocp_nlp_out_set(c->config, c->dims, c->out, c->in, i, "x", v);
        ocp_nlp_out_get(c->config, c->dims, c->out, i + 1, "u", v);
        ocp_nlp_out_set(c->config, c->dims, c->out, c->in, i, "u", v);
    }
    ocp_nlp_out_get(c->config, c->dims, c->out, N, "x", v);
    ocp_nlp_out_set(c->config, c->dims, c->out, c->in, N - 1, "x", v);
}

/* Whole closed loop (R/simulator.py:199-241 without sleeps / visualisation). Returns #failed steps. */
int rmpc_ref_closed_loop(const rmpc_problem *prob, const double z0[12], int nsteps, rmpc_trace *tr)
{
    rmpc_ref *c = rmpc_ref_create(prob, z0);
    if (!c) return -1;
    const int N = prob->N;
    double z[NX], zn[NX], u0[NU];
    memcpy(z, z0, sizeof(z));
    if (tr->z) memcpy(tr->z, z, sizeof(z));
    int nfail = 0;
    for (int k = 0; k < nsteps; k++)
    {
        int sqp_iter = 0, qp_iter = 0;
        double res[4] = {0, 0, 0, 0}, cost = 0, tt = 0;
        int status = rmpc_ref_step(c, z, u0, &sqp_iter, &qp_iter, res, &cost, &tt);
        nfail += status != 0;
        ur10_plant_rk4(prob->model.wcv, prob->dt, z, u0, zn);
        memcpy(z, zn, sizeof(z));
        if (tr->z) memcpy(tr->z + NX * (k + 1), z, sizeof(z));
        if (tr->u) memcpy(tr->u + NU * k, u0, sizeof(u0));
        if (tr->status) tr->status[k] = status;
        if (tr->sqp_iter) tr->sqp_iter[k] = sqp_iter;
        if (tr->qp_iter) tr->qp_iter[k] = qp_iter;
        if (tr->res) memcpy(tr->res + 4 * k, res, sizeof(res));
        if (tr->cost) tr->cost[k] = cost;
        if (tr->time_tot) tr->time_tot[k] = tt;
        if (tr->xtraj || tr->utraj || tr->pitraj)
            rmpc_ref_get_iterate(c, tr->xtraj ? tr->xtraj + (size_t) k * (N + 1) * NX : NULL,
                                 tr->utraj ? tr->utraj + (size_t) k * N * NU : NULL,
                                 tr->pitraj ? tr->pitraj + (size_t) k * N * NX : NULL);
        if (prob->shift) rmpc_ref_shift(c);
    }
    rmpc_ref_destroy(c);
    return nfail;
}

/* Debug/parity hook: HPIPM's per-iteration statistics table of the LAST QP solved
 * (H/ocp_qp/x_ocp_qp_ipm.c:2951-2960; columns: alpha_aff mu_aff sigma alpha_prim alpha_dual mu res_stat res_eq
 * res_ineq res_comp dual_gap obj lq_fact itref_pred itref_corr lin_res x4).  Returns the iteration count. */
int rmpc_ref_last_qp_stat(rmpc_ref *c, double *stat_out, int max_rows, int *stat_m_out)
{
    ocp_nlp_memory *nlp_mem = NULL;
    ocp_nlp_get(c->solver, "nlp_mem", &nlp_mem);
    ocp_qp_xcond_solver_memory *xm = nlp_mem->qp_solver_mem;
    ocp_qp_hpipm_memory *hm = xm->solver_memory;
    struct d_ocp_qp_ipm_ws *ws = hm->hpipm_workspace;
    int stat_m = ws->stat_m, iter = ws->iter;
    if (stat_m_out) *stat_m_out = stat_m;
    int rows = iter + 1 < max_rows ? iter + 1 : max_rows;
    if (stat_out) memcpy(stat_out, ws->stat, sizeof(double) * (size_t) rows * stat_m);
    return iter;
}

/* Debug/parity hook: count, over a closed loop, how many IPM iterations fell back to the LQ factorisation
 * (stat column 12) and the max number of corrector refinement steps (column 14). */
int rmpc_ref_closed_loop_lq_count(const rmpc_problem *prob, const double z0[12], int nsteps, int *n_ipm_iter,
                                  int *n_lq, int *max_itref)
{
    rmpc_ref *c = rmpc_ref_create(prob, z0);
    if (!c) return -1;
    double z[NX], zn[NX], u0[NU];
    memcpy(z, z0, sizeof(z));
    double *stat = malloc(sizeof(double) * 64 * 32);
    *n_ipm_iter = 0; *n_lq = 0; *max_itref = 0;
    for (int k = 0; k < nsteps; k++)
    {
        rmpc_ref_step(c, z, u0, NULL, NULL, NULL, NULL, NULL);
        int sm = 0;
        int it = rmpc_ref_last_qp_stat(c, stat, 64, &sm);
        for (int i = 1; i <= it && i < 64; i++)
        {
            (*n_ipm_iter)++;
            if (stat[sm * i + 12] != 0.0) (*n_lq)++;
            if ((int) stat[sm * i + 14] > *max_itref) *max_itref = (int) stat[sm * i + 14];
        }
        ur10_plant_rk4(prob->model.wcv, prob->dt, z, u0, zn);
        memcpy(z, zn, sizeof(z));
    }
    free(stat);
    rmpc_ref_destroy(c);
    return 0;
}
