/* A plain C99 consumer of include/madmom_b200.h: the SuperFlux features of one int16 signal through
 * mmb_plan_create + mmb_run_host, no Python and no CUDA headers involved.  tests/test_c_abi_example.py compiles it
 * with gcc (CPU suite: the header is valid C and every symbol links) and runs it on the GPU (gpu suite: the numbers
 * it prints equal the ones the ctypes path produces for the same input).
 *
 *   usage: c_abi_example <recipe.bin> <pcm.raw>
 *   recipe.bin: int32 B, int32 start[B], int32 len[B], float weights[sum len], double fft_window[2048]  (written by
 *               the test: the reference's 24 bands/octave table and np.hanning(2048) / 32767)
 *   pcm.raw:    mono little-endian int16, 44.1 kHz
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/madmom_b200.h"

#define CHECK(call)                                                                   \
  do {                                                                                \
    int rc_ = (call);                                                                 \
    if (rc_ != MMB_OK) {                                                              \
      fprintf(stderr, "%s failed: %d (%s)\n", #call, rc_, mmb_last_error());          \
      return 1;                                                                       \
    }                                                                                 \
  } while (0)

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 2;
  int32_t B = 0;
  if (fread(&B, 4, 1, f) != 1 || B <= 0 || B > 4096) return 2;
  int32_t* start = (int32_t*)malloc(4 * (size_t)B);
  int32_t* len = (int32_t*)malloc(4 * (size_t)B);
  if (fread(start, 4, (size_t)B, f) != (size_t)B || fread(len, 4, (size_t)B, f) != (size_t)B) return 2;
  size_t nnz = 0;
  for (int b = 0; b < B; ++b) nnz += (size_t)len[b];
  float* w = (float*)malloc(4 * nnz);
  if (fread(w, 4, nnz, f) != nnz) return 2;
  const int N = 2048;
  const double hop = 441.0;
  double* win = (double*)malloc(8 * (size_t)N);
  if (fread(win, 8, (size_t)N, f) != (size_t)N) return 2;
  fclose(f);

  /* raw little-endian int16 PCM written by the test */
  FILE* g = fopen(argv[2], "rb");
  if (!g) return 2;
  fseek(g, 0, SEEK_END);
  const int64_t n = (int64_t)(ftell(g) / 2);
  fseek(g, 0, SEEK_SET);
  int16_t* pcm = NULL;
  CHECK(mmb_host_alloc((void**)&pcm, n * 2));
  if (fread(pcm, 2, (size_t)n, g) != (size_t)n) return 2;
  fclose(g);

  mmb_config cfg;
  memset(&cfg, 0, sizeof(cfg));   /* fields added by later library versions default to 0 */
  cfg.frame_size = N; cfg.pcm_dtype = MMB_PCM_I16; cfg.hop_size = hop; cfg.origin = 0; cfg.circular_shift = 0;
  cfg.fft_window = win; cfg.num_bands = B; cfg.band_start = start; cfg.band_len = len; cfg.band_weights = w;
  cfg.log_mode = MMB_LOG_LOG10; cfg.mul = 1.0f; cfg.add = 1.0f;
  cfg.diff_frames = 1; cfg.diff_max_bins = 3; cfg.positive_diffs = 1; cfg.chunk_frames = 0; cfg.device = 0;
  mmb_plan* plan = NULL;
  CHECK(mmb_plan_create(&cfg, &plan));
  mmb_plan_info info;
  CHECK(mmb_plan_get_info(plan, &info));

  const int64_t T = mmb_num_frames(n, hop, 0);
  mmb_track unit = {0, n, 0, 0, T, 0};
  float* diff = NULL;
  float* odf = NULL;
  CHECK(mmb_host_alloc((void**)&diff, T * B * 4));
  CHECK(mmb_host_alloc((void**)&odf, T * 4));
  mmb_outputs out;
  out.stft = NULL; out.ld_stft = 0; out.spec = NULL; out.ld_spec = 0; out.filt = NULL; out.ld_filt = 0;
  out.log = NULL; out.ld_log = 0; out.diff = diff; out.ld_diff = B; out.odf = odf; out.lgd = NULL; out.ld_lgd = 0;
  int32_t launches = 0;
  CHECK(mmb_run_host(plan, &unit, 1, pcm, &out, &launches));

  double sum = 0.0, sum_odf = 0.0;
  for (int64_t i = 0; i < T * B; ++i) sum += diff[i];
  for (int64_t i = 0; i < T; ++i) sum_odf += odf[i];
  printf("version %d frames %lld bands %d warps %d launches %d\n", mmb_version(), (long long)T, info.num_bands,
         info.warps, launches);
  printf("sum_diff %.9e sum_odf %.9e odf10 %.9e odf_last %.9e\n", sum, sum_odf, (double)odf[T > 10 ? 10 : 0],
         (double)odf[T - 1]);
  mmb_host_free(diff); mmb_host_free(odf); mmb_host_free(pcm);
  mmb_plan_destroy(plan);
  free(win); free(w); free(len); free(start);
  return 0;
}
