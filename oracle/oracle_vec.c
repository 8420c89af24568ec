/*
 * oracle_vec.c -- the timed CPU baseline's inner loop.  TEST INFRASTRUCTURE ONLY.
 *
 * Same arithmetic as omo_cosine_distance_f32 in oracle.c (pgvector's
 * cosine_distance, reached from every `<=>` site in SURVEY.md section 0), but
 * this translation unit is compiled the way upstream pgvector builds its
 * Makefile target: -O3 -march=native -ftree-vectorize -fassociative-math
 * -fno-signed-zeros -fno-trapping-math.  The compiler may therefore reorder the
 * three float reductions into SIMD lanes, which is what a Postgres backend
 * doing an exact sequential scan actually executes.  It is used for timing
 * (bench.py cpu_baseline, --impl reference); parity is checked against the
 * strictly ordered variants in oracle.c.
 */
#include <math.h>

double omo_cosine_distance_vec(const float *a, const float *b, int dim) {
  float dot = 0.0f, na = 0.0f, nb = 0.0f;
  for (int i = 0; i < dim; i++) {
    dot += a[i] * b[i];
    na += a[i] * a[i];
    nb += b[i] * b[i];
  }
  double s = (double)dot / sqrt((double)na * (double)nb);
  if (s > 1.0) s = 1.0;
  else if (s < -1.0) s = -1.0;
  return 1.0 - s;
}
