# oracle/ref.mk -- build the UNMODIFIED reference numerical core (acados + HPIPM + BLASFEO)
# straight from the sources where they lie under $(REF) (read-only), with plain gcc.
#
# TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is on the product path; see oracle/README.md.
#
# The reference's own CMake build is NOT run.  This file restates, for the one configuration the
# reference project uses, what those CMake files resolve to:
#   * BLASFEO  TARGET=X64_INTEL_HASWELL, LA=HIGH_PERFORMANCE, MF=PANELMAJ, EXT_DEP=ON, BLAS_API=OFF
#              (A/CMakeLists.txt:65 picks X64_AUTOMATIC, which resolves to HASWELL on any AVX2+FMA host;
#               file list = B/CMakeLists.txt:718-769 kernels + :620-640 hp_pm + :474-490 aux)
#   * HPIPM    TARGET=GENERIC, -O2 -DUSE_C99_MATH   (A/CMakeLists.txt:68; H/CMakeLists.txt:158-170,255-300)
#   * acados   every acados/**/*.c minus the optional QP back-ends (AC/CMakeLists.txt:38-70),
#              plus interfaces/acados_c/*.c, -std=c99, no OpenMP (A/CMakeLists.txt:76,136)
#   * optimisation: CMake "Release" (A/CMakeLists.txt:108-112) appends -O3 -DNDEBUG after the
#              per-project -O2, so -O3 -DNDEBUG is what every object is built with
#   * -DACADOS_SILENT (A/CMakeLists.txt:183-187, an option of the reference build) only mutes prints
# The only generated file in the reference build, blasfeo_target.h (B/CMakeLists.txt:458 writes it
# into the source tree), is written into $(OUT)/include instead.
#
# Outputs go ONLY into oracle/_ref/ (git-ignored, travels to the GPU box with the snapshot).
# No reference source is copied into the repository.

REF    ?= /root/reference
A      := $(REF)/acados
B      := $(A)/external/blasfeo
H      := $(A)/external/hpipm
OUT    ?= $(dir $(abspath $(lastword $(MAKEFILE_LIST))))_ref
OBJ    := $(OUT)/obj
CC     ?= gcc

# ---------------------------------------------------------------- BLASFEO
BLASFEO_DEFS := -DTARGET_X64_INTEL_HASWELL -DLA_HIGH_PERFORMANCE -DMF_PANELMAJ -DK_MAX_STACK=300 \
                -DUSE_C99_MATH -DEXT_DEP -DOS_LINUX
BLASFEO_CFLAGS := -O3 -DNDEBUG -fPIC -m64 -mavx -mavx2 -mfma $(BLASFEO_DEFS) -I$(OUT)/include -I$(B)/include -w
BLASFEO_ASFLAGS := $(BLASFEO_DEFS) -DMACRO_LEVEL=1 -I$(OUT)/include -I$(B)/include -fPIC

BLASFEO_SRC := \
  auxiliary/blasfeo_processor_features.c auxiliary/blasfeo_stdlib.c auxiliary/memory.c \
  auxiliary/d_aux_common.c auxiliary/s_aux_common.c \
  auxiliary/d_aux_ext_dep_common.c auxiliary/s_aux_ext_dep_common.c \
  auxiliary/d_aux_ext_dep.c auxiliary/s_aux_ext_dep.c auxiliary/v_aux_ext_dep_lib.c \
  auxiliary/i_aux_ext_dep_lib.c auxiliary/timing.c \
  auxiliary/d_aux_lib4.c auxiliary/s_aux_lib8.c auxiliary/m_aux_lib48.c \
  blasfeo_hp_pm/d_blas1_lib4.c blasfeo_hp_pm/d_blas2_lib4.c blasfeo_hp_pm/d_blas2_diag_lib.c \
  blasfeo_hp_pm/d_blas3_lib4.c blasfeo_hp_pm/d_blas3_diag_lib4.c blasfeo_hp_pm/d_lapack_lib4.c \
  blasfeo_hp_pm/s_blas1_lib8.c blasfeo_hp_pm/s_blas2_lib8.c blasfeo_hp_pm/s_blas2_diag_lib.c \
  blasfeo_hp_pm/s_blas3_lib8.c blasfeo_hp_pm/s_blas3_diag_lib8.c blasfeo_hp_pm/s_lapack_lib8.c \
  kernel/avx2/kernel_dgemm_12x4_lib4.S kernel/avx2/kernel_dgemm_8x8_lib4.S \
  kernel/avx2/kernel_dgemm_8x4_lib4.S kernel/avx2/kernel_dgemm_4x4_lib4.S \
  kernel/avx2/kernel_dgemv_8_lib4.S kernel/avx2/kernel_dgemv_4_lib4.S \
  kernel/avx2/kernel_dsymv_6_lib4.S kernel/avx2/kernel_dger_lib4.S \
  kernel/avx2/kernel_dgetrf_pivot_lib4.c kernel/avx2/kernel_dgebp_lib4.S \
  kernel/avx2/kernel_dgelqf_4_lib4.S kernel/avx2/kernel_dgetr_lib4.c \
  kernel/avx/kernel_dgeqrf_4_lib4.c kernel/avx/kernel_dgemm_diag_lib4.c \
  kernel/avx/kernel_dgecp_lib4.c kernel/avx/kernel_dpack_lib4.S kernel/avx/kernel_dgetr_lib.S \
  kernel/generic/kernel_dgemv_4_lib4.c kernel/generic/kernel_dsymv_4_lib4.c \
  kernel/generic/kernel_dpack_buffer_lib4.c kernel/generic/kernel_dger_lib4.c \
  kernel/generic/kernel_ddot_lib.c kernel/generic/kernel_daxpy_lib.c \
  kernel/avx2/kernel_sgemm_24x4_lib8.S kernel/avx2/kernel_sgemm_16x4_lib8.S \
  kernel/avx2/kernel_sgemm_8x8_lib8.S kernel/avx2/kernel_sgemm_8x4_lib8.S \
  kernel/avx/kernel_sgemm_diag_lib8.c kernel/avx/kernel_sgemv_8_lib8.S \
  kernel/avx/kernel_sgemv_4_lib8.S kernel/avx/kernel_sgecpsc_lib8.S \
  kernel/avx/kernel_sgead_lib8.S kernel/avx/kernel_sgetr_lib8.S kernel/avx/kernel_spack_lib8.S \
  kernel/generic/kernel_sgemm_8x4_lib8.c kernel/generic/kernel_sgemm_4x4_lib4.c \
  kernel/generic/kernel_spack_lib4.c kernel/generic/kernel_sdot_lib.c \
  kernel/generic/kernel_saxpy_lib.c kernel/avx/kernel_d_aux_lib.S \
  kernel/sse3/kernel_align_x64.S
BLASFEO_OBJ := $(patsubst %,$(OBJ)/blasfeo/%.o,$(BLASFEO_SRC))

# ---------------------------------------------------------------- HPIPM (double precision part)
HPIPM_CFLAGS := -O3 -DNDEBUG -fPIC -DTARGET_GENERIC -DUSE_C99_MATH -I$(H)/include -I$(OUT)/include -I$(B)/include -w
HPIPM_SRC := \
  cond/d_cond.c cond/d_cond_aux.c cond/d_part_cond.c cond/d_cond_qcqp.c cond/d_cast_qcqp.c \
  cond/d_part_cond_qcqp.c \
  dense_qp/d_dense_qp.c dense_qp/d_dense_qp_dim.c dense_qp/d_dense_qp_sol.c dense_qp/d_dense_qp_kkt.c \
  dense_qp/d_dense_qp_ipm.c dense_qp/d_dense_qp_res.c dense_qp/d_dense_qp_utils.c \
  dense_qp/d_dense_qp_seed.c dense_qp/d_dense_qcqp_dim.c dense_qp/d_dense_qcqp.c \
  dense_qp/d_dense_qcqp_sol.c dense_qp/d_dense_qcqp_res.c dense_qp/d_dense_qcqp_ipm.c \
  dense_qp/d_dense_qcqp_utils.c \
  ipm_core/d_core_qp_ipm_aux.c ipm_core/d_core_qp_ipm.c \
  ocp_qp/d_ocp_qp.c ocp_qp/d_ocp_qp_dim.c ocp_qp/d_ocp_qp_sol.c ocp_qp/d_ocp_qp_kkt.c \
  ocp_qp/d_ocp_qp_ipm.c ocp_qp/d_ocp_qp_res.c ocp_qp/d_ocp_qp_utils.c ocp_qp/d_ocp_qp_red.c \
  ocp_qp/d_ocp_qp_seed.c ocp_qp/d_ocp_qcqp.c ocp_qp/d_ocp_qcqp_dim.c ocp_qp/d_ocp_qcqp_sol.c \
  ocp_qp/d_ocp_qcqp_ipm.c ocp_qp/d_ocp_qcqp_res.c ocp_qp/d_ocp_qcqp_utils.c \
  auxiliary/aux_string.c auxiliary/aux_mem.c auxiliary/timing.c
HPIPM_OBJ := $(patsubst %,$(OBJ)/hpipm/%.o,$(HPIPM_SRC))

# ---------------------------------------------------------------- acados core + C interface
ACADOS_CFLAGS := -O3 -DNDEBUG -fPIC -std=c99 -DACADOS_SILENT -I$(A) -I$(A)/interfaces -I$(A)/external \
                 -I$(H)/include -I$(OUT)/include -I$(B)/include -w
ACADOS_EXCLUDE := ocp_qp/ocp_qp_hpmpc.c ocp_qp/ocp_qp_ooqp.c dense_qp/dense_qp_ooqp.c \
                  dense_qp/dense_qp_qore.c dense_qp/dense_qp_qpoases.c dense_qp/dense_qp_daqp.c \
                  ocp_qp/ocp_qp_qpdunes.c ocp_qp/ocp_qp_osqp.c ocp_qp/ocp_qp_clarabel.c
ACADOS_SRC := $(filter-out $(ACADOS_EXCLUDE),$(patsubst $(A)/acados/%,%,\
                $(wildcard $(A)/acados/utils/*.c $(A)/acados/ocp_nlp/*.c $(A)/acados/ocp_qp/*.c \
                           $(A)/acados/dense_qp/*.c $(A)/acados/sim/*.c)))
ACADOS_OBJ := $(patsubst %,$(OBJ)/acados/%.o,$(ACADOS_SRC))
ACADOSC_SRC := $(notdir $(wildcard $(A)/interfaces/acados_c/*.c))
ACADOSC_OBJ := $(patsubst %,$(OBJ)/acados_c/%.o,$(ACADOSC_SRC))

.PHONY: ref clean-ref
ref: $(OUT)/libacados_ref.a

$(OUT)/include/blasfeo_target.h:
	@mkdir -p $(dir $@)
	@printf '%s\n' '#ifndef TARGET_X64_INTEL_HASWELL' '#define TARGET_X64_INTEL_HASWELL' '#endif' \
	  '#ifndef TARGET_NEED_FEATURE_AVX2' '#define TARGET_NEED_FEATURE_AVX2 1' '#endif' \
	  '#ifndef TARGET_NEED_FEATURE_FMA' '#define TARGET_NEED_FEATURE_FMA 1' '#endif' \
	  '#ifndef LA_HIGH_PERFORMANCE' '#define LA_HIGH_PERFORMANCE' '#endif' \
	  '#ifndef MF_PANELMAJ' '#define MF_PANELMAJ' '#endif' \
	  '#ifndef EXT_DEP' '#define EXT_DEP' '#endif' > $@

$(OBJ)/blasfeo/%.c.o: $(B)/%.c $(OUT)/include/blasfeo_target.h
	@mkdir -p $(dir $@)
	@$(CC) $(BLASFEO_CFLAGS) -c $< -o $@
$(OBJ)/blasfeo/%.S.o: $(B)/%.S $(OUT)/include/blasfeo_target.h
	@mkdir -p $(dir $@)
	@$(CC) $(BLASFEO_ASFLAGS) -c $< -o $@
$(OBJ)/hpipm/%.c.o: $(H)/%.c $(OUT)/include/blasfeo_target.h
	@mkdir -p $(dir $@)
	@$(CC) $(HPIPM_CFLAGS) -c $< -o $@
$(OBJ)/acados/%.c.o: $(A)/acados/%.c $(OUT)/include/blasfeo_target.h
	@mkdir -p $(dir $@)
	@$(CC) $(ACADOS_CFLAGS) -c $< -o $@
$(OBJ)/acados_c/%.c.o: $(A)/interfaces/acados_c/%.c $(OUT)/include/blasfeo_target.h
	@mkdir -p $(dir $@)
	@$(CC) $(ACADOS_CFLAGS) -c $< -o $@

$(OUT)/libacados_ref.a: $(BLASFEO_OBJ) $(HPIPM_OBJ) $(ACADOS_OBJ) $(ACADOSC_OBJ)
	@rm -f $@
	@ar rcs $@ $^ && echo "built $@"

clean-ref:
	rm -rf $(OUT)

# ---------------------------------------------------------------- driver (our code) + reference archive -> one .so
DRIVER_CFLAGS := -O2 -fPIC -std=c99 -I$(A) -I$(A)/interfaces -I$(A)/external -I$(H)/include \
                 -I$(OUT)/include -I$(B)/include -I$(dir $(abspath $(lastword $(MAKEFILE_LIST))))
ORACLE_DIR := $(dir $(abspath $(lastword $(MAKEFILE_LIST))))

.PHONY: ref-driver
ref-driver: $(OUT)/librmpc_ref.so

$(OUT)/librmpc_ref.so: $(ORACLE_DIR)ref_driver.c $(ORACLE_DIR)ur10_model.c $(ORACLE_DIR)ur10_model.h \
                       $(ORACLE_DIR)rmpc_problem.h $(OUT)/libacados_ref.a
	$(CC) $(DRIVER_CFLAGS) -shared -o $@ $(ORACLE_DIR)ref_driver.c $(ORACLE_DIR)ur10_model.c \
	    -Wl,--whole-archive $(OUT)/libacados_ref.a -Wl,--no-whole-archive -lm
