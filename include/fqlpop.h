/* fqlpop.h -- C-ABI of the B200-native population FQL hot path (libfqlpop.so).
 *
 * The reference (Mouafak-Khaled/Flow-Q-Learning) has no FFI of its own: its hot path is reached through
 * Python duck-typing (FQLAgent.update at trainer/experiment.py:109, FQLAgent.sample_actions at
 * evaluator/evaluation.py:58-63, OfflineTaskWithSimulatedEvaluations.step at
 * task/offline_task_simulated.py:85-107, Dataset.sample at task/offline_task_simulated.py:53-58).
 * This header is the boundary a maintainer binds with ctypes to replace those calls (INTEGRATION.md shows
 * the stub). Plain pointers and sizes only; every function returns 0 on success and a non-zero status
 * otherwise, with a message available from fqlpop_last_error(). No exceptions cross the ABI.
 * Device pointers are marked (dev); all others are host pointers. `stream` is a cudaStream_t (may be NULL).
 */
#ifndef FQLPOP_H_
#define FQLPOP_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* fqlpop_last_error(void);
int fqlpop_version(void);

/* ---- kernel-level entry: one grouped GEMM with a fused epilogue (unit tests / micro-benchmarks) ------
 * Replaces a single nn.Dense (+GELU(+LayerNorm)) or its transpose inside FQLAgent.update.            */
typedef struct {
  int32_t M, N, K;        /* rows, cols (multiple of 64), reduction (multiple of 64)                   */
  int32_t P;              /* batch = population members                                                */
  int32_t a_mn, b_mn;     /* 0: operand is K-major ([rows][K]); 1: MN-major ([K][rows])                */
  int32_t n_sub;          /* 1: 128x128 tiles; 4: 128x512 tiles (required by the LayerNorm epilogues)  */
  int32_t epi_kind, act;  /* EpiKind / ActKind (see csrc/gemm.h)                                       */
  float ln_eps;
  const void* A; int64_t a_ld, a_ms;   /* (dev) bf16 */
  const void* B; int64_t b_ld, b_ms;   /* (dev) bf16 */
  const float* bias; const float* gamma; const float* beta;          /* (dev) [P][N]                   */
  void* out_h; void* out_gp; void* out_xhat; float* out_rstd;        /* (dev) bf16 [P][M][N]; [P][M]   */
  const void* gp; const void* xhat; const float* rstd;               /* (dev)                          */
  float* out_f32;                                                    /* (dev) [P][M][N]                */
  float* dbias_part; float* dgamma_part; float* dbeta_part;          /* (dev) [P][ceil(M/128)][N]      */
  const int32_t* n_active; const int32_t* member_list;               /* (dev) optional                 */
} fqlpop_gemm_test_t;
int fqlpop_gemm_test(const fqlpop_gemm_test_t* desc, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FQLPOP_H_ */
