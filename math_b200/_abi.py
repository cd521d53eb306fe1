"""ctypes mirror of include/stan_cuda.h (structs, flags, prototypes).

Pure declarations: importing this module loads nothing.  ``bind(lib)`` attaches the prototypes
to a ``ctypes.CDLL`` of libstanmath_cuda.so; the test-side oracle loaders reuse the structs so
that identical argument blocks can be handed to the CUDA library and to the CPU checkers.
"""
import ctypes as C

OK, NONFINITE, BAD_OUTCOME, BAD_OPERAND, LOG_ZERO, NOT_POS_DEF, NOT_SYMMETRIC = 0, 1, 2, 3, 4, 5, 6
ERR_ARG, ERR_CUDA, ERR_NCCL, ERR_STATE = -1, -2, -3, -4

BERNOULLI_LOGIT, POISSON_LOG, NORMAL_ID, NEG_BINOMIAL_2_LOG, ORDERED_LOGISTIC = 0, 1, 2, 3, 4
CATEGORICAL_LOGIT = 5
FAMILIES = {"bernoulli_logit": BERNOULLI_LOGIT, "poisson_log": POISSON_LOG, "normal_id": NORMAL_ID,
            "neg_binomial_2_log": NEG_BINOMIAL_2_LOG, "ordered_logistic": ORDERED_LOGISTIC,
            "categorical_logit": CATEGORICAL_LOGIT}
HAS_SCALE = (NORMAL_ID, NEG_BINOMIAL_2_LOG)   # families with a fifth operand (sigma | phi)

LP_CONST, LP_SCALE, LP_MAIN, LP_ALL = 0x01, 0x02, 0x04, 0x07
LP_LINEAR = 0x200
NEED_DBETA, NEED_DALPHA, NEED_DSIGMA, NEED_DY, NEED_ALL = 0x08, 0x10, 0x20, 0x40, 0x78
NEED_DCUTS = 0x20
NEED_DX = 0x400
LP_DOF = 0x800      # student_t_lpdf: lgamma(nu/2 + 1/2) - lgamma(nu/2) - log(nu)/2
NEED_DNU = 0x08     # student_t_lpdf: d/dnu (shares the bit of NEED_DBETA)
FUSED_ALLREDUCE = 0x100

c_double_p = C.POINTER(C.c_double)


class GlmArgs(C.Structure):
    _fields_ = [
        ("family", C.c_int32),
        ("flags", C.c_uint32),
        ("n", C.c_int64),
        ("k", C.c_int64),
        ("x", C.c_void_p),
        ("ldx", C.c_int64),
        ("x_row_major", C.c_int32),
        ("y_is_double", C.c_int32),
        ("y", C.c_void_p),
        ("alpha", C.c_void_p),
        ("alpha_len", C.c_int64),
        ("beta", C.c_void_p),
        ("sigma", C.c_void_p),
        ("sigma_len", C.c_int64),
        ("lp", C.c_void_p),
        ("dbeta", C.c_void_p),
        ("dalpha", C.c_void_p),
        ("dsigma", C.c_void_p),
        ("dy", C.c_void_p),
        ("aux", C.c_void_p),
        ("cuts", C.c_void_p),
        ("cuts_len", C.c_int64),
        ("dcuts", C.c_void_p),
        ("n_classes", C.c_int64),
        ("y_per_call", C.c_int32),
        ("x_per_call", C.c_int32),
        ("dx", C.c_void_p),
        ("lddx", C.c_int64),
    ]


class NormalArgs(C.Structure):
    _fields_ = [
        ("flags", C.c_uint32),
        ("y_is_data", C.c_int32),
        ("n", C.c_int64),
        ("y", C.c_void_p),
        ("y_len", C.c_int64),
        ("mu", C.c_void_p),
        ("mu_len", C.c_int64),
        ("sigma", C.c_void_p),
        ("sigma_len", C.c_int64),
        ("lp", C.c_void_p),
        ("dmu", C.c_void_p),
        ("dsigma", C.c_void_p),
        ("dy", C.c_void_p),
    ]


class StudentTArgs(C.Structure):
    _fields_ = [
        ("flags", C.c_uint32),
        ("y_is_data", C.c_int32),
        ("n", C.c_int64),
        ("y", C.c_void_p),
        ("y_len", C.c_int64),
        ("nu", C.c_void_p),
        ("nu_len", C.c_int64),
        ("mu", C.c_void_p),
        ("mu_len", C.c_int64),
        ("sigma", C.c_void_p),
        ("sigma_len", C.c_int64),
        ("lp", C.c_void_p),
        ("dy", C.c_void_p),
        ("dnu", C.c_void_p),
        ("dmu", C.c_void_p),
        ("dsigma", C.c_void_p),
    ]


class MultiplyArgs(C.Structure):
    _fields_ = [
        ("m", C.c_int64),
        ("n", C.c_int64),
        ("k", C.c_int64),
        ("a", C.c_void_p),
        ("lda", C.c_int64),
        ("trans_a", C.c_int32),
        ("trans_b", C.c_int32),
        ("b", C.c_void_p),
        ("ldb", C.c_int64),
        ("c", C.c_void_p),
        ("ldc", C.c_int64),
    ]


(LOCSCALE_NORMAL, LOCSCALE_CAUCHY, LOCSCALE_LOGNORMAL, LOCSCALE_GAMMA, LOCSCALE_EXPONENTIAL, LOCSCALE_BETA, LOCSCALE_WEIBULL,
 LOCSCALE_INV_GAMMA) = range(8)
LOCSCALES = {"normal": LOCSCALE_NORMAL, "cauchy": LOCSCALE_CAUCHY, "lognormal": LOCSCALE_LOGNORMAL, "gamma": LOCSCALE_GAMMA,
             "exponential": LOCSCALE_EXPONENTIAL, "beta": LOCSCALE_BETA, "weibull": LOCSCALE_WEIBULL, "inv_gamma": LOCSCALE_INV_GAMMA}
LP_Y = 0x1000
DENSITY_BERNOULLI_LOGIT, DENSITY_POISSON_LOG, DENSITY_POISSON, DENSITY_BERNOULLI = 1, 2, 3, 4
DENSITIES = {"bernoulli_logit": DENSITY_BERNOULLI_LOGIT, "poisson_log": DENSITY_POISSON_LOG, "poisson": DENSITY_POISSON,
             "bernoulli": DENSITY_BERNOULLI}


class DensityArgs(C.Structure):
    _fields_ = [
        ("density", C.c_int32),
        ("flags", C.c_uint32),
        ("n", C.c_int64),
        ("y", C.c_void_p),
        ("y_len", C.c_int64),
        ("theta", C.c_void_p),
        ("theta_len", C.c_int64),
        ("lp", C.c_void_p),
        ("dtheta", C.c_void_p),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("cache_hits", C.c_uint64),
        ("cache_misses", C.c_uint64),
        ("bytes_uploaded", C.c_uint64),
        ("evals", C.c_uint64),
        ("kernel_launches", C.c_uint64),
        ("allreduces", C.c_uint64),
        ("last_fill_ms", C.c_double),
        ("fused_allreduces", C.c_uint64),
        ("cache_evictions", C.c_uint64),
    ]


# name -> (restype, argtypes); exactly the symbols include/stan_cuda.h declares
PROTOTYPES = {
    "stan_cuda_init": (C.c_int, [C.c_int]),
    "stan_cuda_init_device": (C.c_int, [C.c_int]),
    "stan_cuda_device_count": (C.c_int, []),
    "stan_cuda_describe": (C.c_char_p, []),
    "stan_cuda_last_error": (C.c_char_p, []),
    "stan_cuda_stats": (C.c_int, [C.POINTER(Stats)]),
    "stan_cuda_stats_reset": (None, []),
    "stan_cuda_set_shard_min_elems": (None, [C.c_int64]),
    "stan_cuda_set_tuning": (None, [C.c_int, C.c_int, C.c_int]),
    "stan_cuda_set_kernel": (None, [C.c_int]),
    "stan_cuda_set_categorical_kernel": (None, [C.c_int]),
    "stan_cuda_glm_eval": (C.c_int, [C.POINTER(GlmArgs)]),
    "stan_cuda_glm_eval_batch": (C.c_int, [C.POINTER(GlmArgs), C.c_int, C.c_int, C.POINTER(C.c_int)]),
    "stan_cuda_density_eval": (C.c_int, [C.POINTER(DensityArgs)]),
    "stan_cuda_locscale_lpdf_eval": (C.c_int, [C.c_int32, C.POINTER(NormalArgs)]),
    "stan_cuda_normal_lpdf_eval": (C.c_int, [C.POINTER(NormalArgs)]),
    "stan_cuda_student_t_lpdf_eval": (C.c_int, [C.POINTER(StudentTArgs)]),
    "stan_cuda_matrix_create": (C.c_int, [C.c_int64, C.c_int64, C.POINTER(C.c_void_p)]),
    "stan_cuda_matrix_free": (None, [C.c_void_p]),
    "stan_cuda_matrix_rows": (C.c_int64, [C.c_void_p]),
    "stan_cuda_matrix_cols": (C.c_int64, [C.c_void_p]),
    "stan_cuda_matrix_upload": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "stan_cuda_matrix_download": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "stan_cuda_matrix_multiply": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]),
    "stan_cuda_multiply": (C.c_int, [C.POINTER(MultiplyArgs)]),
    "stan_cuda_cholesky_decompose": (C.c_int, [C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_double]),
    "stan_cuda_gp_exp_quad_cov": (C.c_int, [C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_double, C.c_double,
                                            C.c_void_p, C.c_int64]),
    "stan_cuda_tri_solve": (C.c_int, [C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                      C.c_void_p, C.c_int64]),
    "stan_cuda_cholesky_adjoint": (C.c_int, [C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]),
    "stan_cuda_host_staging": (C.c_void_p, [C.c_size_t]),
    "stan_cuda_cache_invalidate": (C.c_int, [C.c_void_p]),
    "stan_cuda_cache_clear": (None, []),
    "stan_cuda_cache_set_limit": (None, [C.c_int64]),
    "stan_cuda_cache_bytes": (C.c_int64, []),
    "stan_cuda_cache_set_verify": (None, [C.c_int]),
    "stan_cuda_shard_scope_enter": (None, [C.c_int, C.c_int, C.c_int]),
    "stan_cuda_shard_scope_exit": (None, []),
    "stan_cuda_comm_unique_id": (C.c_int, [C.c_void_p]),
    "stan_cuda_comm_init_rank": (C.c_int, [C.c_int, C.c_int, C.c_void_p]),
    "stan_cuda_comm_world_size": (C.c_int, []),
    "stan_cuda_comm_fused_available": (C.c_int, []),
    "stan_cuda_comm_describe": (C.c_char_p, []),
    "stan_cuda_comm_destroy": (None, []),
    "stan_cuda_data_create": (C.c_int, [C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_int64,
                                        C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]),
    "stan_cuda_data_destroy": (None, [C.c_void_p]),
    "stan_cuda_data_params_len": (C.c_int64, [C.c_void_p]),
    "stan_cuda_data_result_len": (C.c_int64, [C.c_void_p]),
    "stan_cuda_data_kpad": (C.c_int64, [C.c_void_p]),
    "stan_cuda_data_algorithmic_bytes": (C.c_int64, [C.c_void_p]),
    "stan_cuda_glm_launch": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "stan_cuda_glm_launch_ordered": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "stan_cuda_glm_launch_categorical": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "stan_cuda_glm_launch_noalloc": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "stan_cuda_allreduce": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p]),
    "stan_cuda_data_launches_per_eval": (C.c_int, [C.c_void_p]),
    "stan_cuda_data_y_range": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
}


def bind(lib):
    """Attach restype/argtypes for every exported entry point; raises if one is missing."""
    for name, (restype, argtypes) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = restype
        fn.argtypes = argtypes
    return lib
