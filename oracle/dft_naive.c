/*
 * dft_naive.c — TEST INFRASTRUCTURE ONLY. The reference's DFT interface (lib/include/srslte/phy/dft/dft.h:76-158) on a
 * direct O(N^2) transform, so that the reference's own test programs (pusch_test: SC-FDMA transform precoding,
 * dft_precoding.c:104-118) link and run in this image, which has no FFTW (lib/src/phy/dft/dft_fftw.c needs it).
 * Complex transforms only (plain, batched "guru" with strides, and the mirror / dc / norm / db options of srslte_dft_run_c,
 * dft_fftw.c:262-309); the real-to-real plans are not provided. Double-precision accumulation over a twiddle table.
 */
#include <complex.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "srslte/phy/dft/dft.h"

typedef struct {
  int n, sign;
  double complex *tw;                       /* exp(sign * 2 pi i k / n) */
  cf_t *gin, *gout;                         /* guru plans: caller buffers and strides */
  int istride, ostride, how_many, idist, odist;
} naive_plan_t;

static naive_plan_t *np_new(int n, int sign)
{
  naive_plan_t *p = calloc(1, sizeof(*p));
  if (!p) return NULL;
  p->n = n; p->sign = sign;
  p->tw = malloc(sizeof(double complex) * (size_t)(n > 0 ? n : 1));
  for (int k = 0; k < n; k++) p->tw[k] = cexp(sign * 2.0 * M_PI * I * (double)k / (double)n);
  p->istride = p->ostride = 1; p->how_many = 1;
  return p;
}
static void np_free(naive_plan_t *p) { if (p) { free(p->tw); free(p); } }

static void np_exec(const naive_plan_t *p, const cf_t *in, int is, cf_t *out, int os)
{
  const int n = p->n;
  cf_t *tmp = malloc(sizeof(cf_t) * (size_t)n);          /* in may alias out */
  for (int k = 0; k < n; k++) {
    double complex acc = 0;
    long idx = 0;
    for (int j = 0; j < n; j++) { acc += (double complex)in[(size_t)j * is] * p->tw[idx]; idx += k; if (idx >= n) idx -= n; }
    tmp[k] = (cf_t)acc;
  }
  for (int k = 0; k < n; k++) out[(size_t)k * os] = tmp[k];
  free(tmp);
}

void srslte_dft_load(void) { }
void srslte_dft_exit(void) { }

static void plan_defaults(srslte_dft_plan_t *plan, int n, srslte_dft_dir_t dir)
{
  plan->size = plan->init_size = n;
  plan->mode = SRSLTE_DFT_COMPLEX; plan->dir = dir; plan->forward = dir == SRSLTE_DFT_FORWARD;
  plan->mirror = plan->db = plan->norm = plan->dc = false;
}

int srslte_dft_plan_c(srslte_dft_plan_t *plan, const int dft_points, srslte_dft_dir_t dir)
{
  plan->in = malloc(sizeof(cf_t) * (size_t)dft_points);
  plan->out = malloc(sizeof(cf_t) * (size_t)dft_points);
  plan->p = np_new(dft_points, dir == SRSLTE_DFT_FORWARD ? -1 : 1);
  if (!plan->in || !plan->out || !plan->p) return -1;
  plan_defaults(plan, dft_points, dir);
  plan->is_guru = false;
  return 0;
}

int srslte_dft_plan_guru_c(srslte_dft_plan_t *plan, const int dft_points, srslte_dft_dir_t dir, cf_t *in_buffer, cf_t *out_buffer,
                           int istride, int ostride, int how_many, int idist, int odist)
{
  naive_plan_t *p = np_new(dft_points, dir == SRSLTE_DFT_FORWARD ? -1 : 1);
  if (!p) return -1;
  p->gin = in_buffer; p->gout = out_buffer; p->istride = istride; p->ostride = ostride;
  p->how_many = how_many; p->idist = idist; p->odist = odist;
  plan->p = p;
  plan_defaults(plan, dft_points, dir);
  plan->is_guru = true;
  return 0;
}

int srslte_dft_plan_r(srslte_dft_plan_t *plan, const int dft_points, srslte_dft_dir_t dir)
{
  (void)plan; (void)dft_points; (void)dir;
  fprintf(stderr, "dft_naive: real-to-real plans are not provided\n");
  return -1;
}

int srslte_dft_plan(srslte_dft_plan_t *plan, const int dft_points, srslte_dft_dir_t dir, srslte_dft_mode_t mode)
{
  memset(plan, 0, sizeof(*plan));
  return mode == SRSLTE_DFT_COMPLEX ? srslte_dft_plan_c(plan, dft_points, dir) : srslte_dft_plan_r(plan, dft_points, dir);
}

int srslte_dft_replan_c(srslte_dft_plan_t *plan, const int new_dft_points)
{
  np_free(plan->p);
  plan->p = np_new(new_dft_points, plan->dir == SRSLTE_DFT_FORWARD ? -1 : 1);
  if (!plan->p) return -1;
  plan->size = new_dft_points;
  return 0;
}

int srslte_dft_replan_guru_c(srslte_dft_plan_t *plan, const int new_dft_points, cf_t *in_buffer, cf_t *out_buffer,
                             int istride, int ostride, int how_many, int idist, int odist)
{
  const srslte_dft_dir_t dir = plan->forward ? SRSLTE_DFT_FORWARD : SRSLTE_DFT_BACKWARD;
  np_free(plan->p);
  return srslte_dft_plan_guru_c(plan, new_dft_points, dir, in_buffer, out_buffer, istride, ostride, how_many, idist, odist);
}

int srslte_dft_replan_r(srslte_dft_plan_t *plan, const int new_dft_points) { (void)plan; (void)new_dft_points; return -1; }

int srslte_dft_replan(srslte_dft_plan_t *plan, const int new_dft_points)
{
  if (new_dft_points > plan->init_size) {
    fprintf(stderr, "DFT: replan to %d points beyond the initial %d\n", new_dft_points, plan->init_size);
    return -1;
  }
  return plan->mode == SRSLTE_DFT_COMPLEX ? srslte_dft_replan_c(plan, new_dft_points) : srslte_dft_replan_r(plan, new_dft_points);
}

void srslte_dft_plan_free(srslte_dft_plan_t *plan)
{
  if (!plan) return;
  if (!plan->is_guru) { free(plan->in); free(plan->out); }
  np_free(plan->p);
  memset(plan, 0, sizeof(*plan));
}

void srslte_dft_plan_set_mirror(srslte_dft_plan_t *plan, bool val) { plan->mirror = val; }
void srslte_dft_plan_set_db(srslte_dft_plan_t *plan, bool val) { plan->db = val; }
void srslte_dft_plan_set_norm(srslte_dft_plan_t *plan, bool val) { plan->norm = val; }
void srslte_dft_plan_set_dc(srslte_dft_plan_t *plan, bool val) { plan->dc = val; }

void srslte_dft_run_c_zerocopy(srslte_dft_plan_t *plan, const cf_t *in, cf_t *out) { np_exec(plan->p, in, 1, out, 1); }

void srslte_dft_run_c(srslte_dft_plan_t *plan, const cf_t *in, cf_t *out)
{
  const int n = plan->size, off = plan->dc ? 1 : 0;
  cf_t *pi = plan->in, *po = plan->out;
  /* input side of the "mirror" option (negative and positive frequencies swapped, DC carrier left out): inverse transforms */
  if (plan->mirror && !plan->forward) {
    const int h = n / 2;
    for (int i = 0; i < off; i++) pi[i] = 0;
    memcpy(pi + off, in + h, sizeof(cf_t) * (size_t)(n - h - off));
    memcpy(pi + (n - h), in, sizeof(cf_t) * (size_t)h);
  } else {
    memcpy(pi, in, sizeof(cf_t) * (size_t)n);
  }
  np_exec(plan->p, pi, 1, po, 1);
  if (plan->norm) { const float s = 1.0f / sqrtf((float)n); for (int i = 0; i < n; i++) po[i] *= s; }
  if (plan->db) for (int i = 0; i < n; i++) po[i] = 10 * log10(po[i]);
  /* output side of it: forward transforms */
  if (plan->mirror && plan->forward) {
    const int h = (n + 1) / 2;
    memcpy(out, po + h, sizeof(cf_t) * (size_t)(n - h));
    memcpy(out + (n - h), po + off, sizeof(cf_t) * (size_t)(h - off));
  } else {
    memcpy(out, po, sizeof(cf_t) * (size_t)n);
  }
}

void srslte_dft_run_guru_c(srslte_dft_plan_t *plan)
{
  if (!plan->is_guru) { fprintf(stderr, "srslte_dft_run_guru_c: the selected plan is not guru!\n"); return; }
  const naive_plan_t *p = plan->p;
  for (int b = 0; b < p->how_many; b++)
    np_exec(p, p->gin + (size_t)b * p->idist, p->istride, p->gout + (size_t)b * p->odist, p->ostride);
}

void srslte_dft_run_r(srslte_dft_plan_t *plan, const float *in, float *out) { (void)plan; (void)in; (void)out; }

void srslte_dft_run(srslte_dft_plan_t *plan, const void *in, void *out)
{
  if (plan->mode == SRSLTE_DFT_COMPLEX) srslte_dft_run_c(plan, in, out);
}
