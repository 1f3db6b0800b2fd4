/*
 * b2dp_oracle.c -- CPU restatement of the reference's data-parallel hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may build, load or call it, and only as the checker or as the timed CPU
 * baseline.  The shipped library (libb200dp.so) never links or calls this file.
 *
 * PARITY UNPINNED.  The reference (/root/reference, 8 Keras scripts) holds no
 * implementation of this arithmetic, no tests and no golden vectors; the sums
 * and the optimizer live in TensorFlow/Keras and Horovod, which are not vendored,
 * not version-pinned and not installable on this image.  Every function below
 * therefore restates the *published* upstream algorithm from memory of the public
 * source trees ([upstream]) and cites the reference call site that reaches it.
 * What pins it instead (tests/test_oracle.py): an independent numpy float32
 * restatement (oracle/oracle_np.py), torch's bf16 cast, torch.optim.Adam/SGD
 * under a stated tolerance, and committed golden vectors under tests/golden/.
 *
 * Arithmetic contract: IEEE-754 binary32, round-to-nearest-even after EVERY
 * operation, no fused multiply-add (build with -ffp-contract=off; the Makefile
 * does).  That is what TensorFlow's CPU kernels (Eigen, scalar/packet add, mul,
 * sub, div, sqrt) give for the reference's own CPU-runnable config
 * (imagenet-resnet50.py, BASELINE.json configs[0]).
 *
 * OpenMP is used only to time the baseline on all host cores; every loop is
 * elementwise so the thread count cannot change a single bit of the result.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

int orc_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void orc_set_threads(int t) {
#ifdef _OPENMP
  if (t > 0) omp_set_num_threads(t);
#else
  (void)t;
#endif
}

/* ------------------------------------------------------------------------- */
/* bf16 wire format: round-to-nearest-even truncation of binary32, NaN kept quiet.
 * This is the cast a "bf16 gradient mode" performs before the wire
 * (BASELINE.json north_star; SURVEY.md 8c "bf16 gradient mode").                */
uint16_t orc_f32_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x0040u); /* NaN */
  uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;
  return (uint16_t)(u >> 16);
}

float orc_bf16_to_f32(uint16_t h) {
  uint32_t u = ((uint32_t)h) << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

void orc_cast_bf16(const float* x, uint16_t* out, size_t len) {
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < len; ++i) out[i] = orc_f32_to_bf16(x[i]);
}

/* ------------------------------------------------------------------------- */
/* Reduction.
 *
 * MirroredStrategy() (imagenet-resnet50-mirror.py:21) reduces each step's
 * gradients under model.fit (:81) with batch_reduce(ReduceOp.SUM, ...).  The
 * deterministic form of that reduction is tf.distribute.ReductionToOneDevice:
 * [upstream tensorflow/python/distribute/cross_device_ops.py::_simple_reduce ->
 * cross_device_utils.aggregate_tensors_or_indexed_slices -> math_ops.add_n ->
 * core/kernels/aggregate_ops.cc] = ((((g0 + g1) + g2) + ...) + g_{n-1}) in
 * fp32, replica order, one rounding per add.  prescale is applied to each input
 * before the add (Horovod's prescale_factor; 1.0 for tf.distribute).            */
void orc_addn_f32(const float* const* g, int n, size_t len, float prescale, float* out) {
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < len; ++i) {
    float acc = (prescale == 1.0f) ? g[0][i] : g[0][i] * prescale;
    for (int r = 1; r < n; ++r) {
      float x = (prescale == 1.0f) ? g[r][i] : g[r][i] * prescale;
      acc = acc + x;
    }
    out[i] = acc;
  }
}

/* Same sum with every input first rounded to bf16 (the wire format) and the
 * accumulation kept in fp32, replica order.                                    */
void orc_addn_bf16wire(const float* const* g, int n, size_t len, float prescale, float* out) {
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < len; ++i) {
    float acc = 0.0f;
    for (int r = 0; r < n; ++r) {
      float x = (prescale == 1.0f) ? g[r][i] : g[r][i] * prescale;
      float q = orc_bf16_to_f32(orc_f32_to_bf16(x));
      acc = (r == 0) ? q : acc + q;
    }
    out[i] = acc;
  }
}

/* ReduceOp.MEAN [upstream cross_device_ops.py::_simple_reduce ->
 * cross_device_utils.divide_by_n_tensors_or_indexed_slices]: the AddN sum
 * divided by n (true division).  hvd.allreduce(op=Average), reached from
 * hvd.DistributedOptimizer (imagenet-resnet50-hvd.py:101), is instead the sum
 * times float(1.0/n) [upstream horovod/common/ops/nccl_operations.cc +
 * cuda/cuda_kernels.cu::scale_buffer_k]; both are exposed as a postscale.      */
void orc_div_n_f32(float* x, size_t len, int n) {
  float fn = (float)n;
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < len; ++i) x[i] = x[i] / fn;
}

void orc_scale_f32(float* x, size_t len, float s) {
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < len; ++i) x[i] = x[i] * s;
}

/* ------------------------------------------------------------------------- */
/* Adam, TensorFlow's fused form.
 *
 * compile(optimizer='adam') (imagenet-resnet50-mirror.py:76, imagenet-resnet50.py:62,
 * -multiworkers.py:90, -ps.py:137) and Adam(learning_rate=0.1*hvd.size())
 * (imagenet-resnet50-hvd.py:99).  [upstream core/kernels/training_ops.cc
 * ApplyAdam<CPUDevice>; keras/optimizers adam.py update_step]:
 *     alpha = lr * sqrt(1 - beta2^t) / (1 - beta1^t)          (fp32 scalars)
 *     m    += (g - m) * (1 - beta1)
 *     v    += (g*g - v) * (1 - beta2)
 *     w    -= (alpha * m) / (sqrt(v) + eps)
 * t is the 1-based step.  lr is an argument because the scripts' callbacks
 * change it while training (ReduceLROnPlateau -mirror.py:78,
 * LearningRateWarmupCallback -hvd.py:115).                                      */
float orc_adam_alpha(float lr, float beta1, float beta2, long step) {
  float b1p = powf(beta1, (float)step);
  float b2p = powf(beta2, (float)step);
  float num = sqrtf(1.0f - b2p);
  float den = 1.0f - b1p;
  float a = lr * num;
  return a / den;
}

void orc_adam_f32(float* w, float* m, float* v, const float* g, size_t len, float lr,
                  float beta1, float beta2, float eps, long step) {
  const float alpha = orc_adam_alpha(lr, beta1, beta2, step);
  const float omb1 = 1.0f - beta1;
  const float omb2 = 1.0f - beta2;
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < len; ++i) {
    float gi = g[i];
    float d1 = gi - m[i];
    float p1 = d1 * omb1;
    float mi = m[i] + p1;
    float g2 = gi * gi;
    float d2 = g2 - v[i];
    float p2 = d2 * omb2;
    float vi = v[i] + p2;
    float num = alpha * mi;
    float den = sqrtf(vi) + eps;
    float upd = num / den;
    m[i] = mi;
    v[i] = vi;
    w[i] = w[i] - upd;
  }
}

/* SGD with momentum, Keras form.  Not used by any reference script (they all
 * use Adam); named by BASELINE.json north_star as kernel (3).
 * [upstream keras/optimizers sgd.py update_step]:
 *     if wd != 0:  w -= (w * wd) * lr          (decoupled weight decay, first)
 *     a  = mu * a - lr * g
 *     w += a                      (nesterov:  w += mu * a - lr * g)
 * mu == 0 degenerates to w -= lr * g with a untouched.                          */
void orc_sgdm_f32(float* w, float* a, const float* g, size_t len, float lr, float mu,
                  int nesterov, float wd) {
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < len; ++i) {
    float wi = w[i];
    float gi = g[i];
    if (wd != 0.0f) {
      float t = wi * wd;
      t = t * lr;
      wi = wi - t;
    }
    float lg = lr * gi;
    if (mu != 0.0f) {
      float ma = mu * a[i];
      float ai = ma - lg;
      a[i] = ai;
      if (nesterov) {
        float t = mu * ai;
        t = t - lg;
        wi = wi + t;
      } else {
        wi = wi + ai;
      }
    } else {
      wi = wi - lg;
    }
    w[i] = wi;
  }
}

/* Sum of squares of the reduced gradient, accumulated in double in index order
 * (used only to check the kernels' optional gradient-norm statistic, which is
 * compared under a relative tolerance, never bitwise).                          */
double orc_sumsq_f64(const float* g, size_t len) {
  double s = 0.0;
  for (size_t i = 0; i < len; ++i) s += (double)g[i] * (double)g[i];
  return s;
}

/* One whole data-parallel step as the reference performs it on n replicas:
 * reduce (SUM, optional bf16 wire, optional postscale), then Adam on the single
 * logical copy of the weights that every replica mirrors.  g holds n replica
 * gradients; gsum (len floats) is scratch/out for the reduced gradient.         */
void orc_step_adam(const float* const* g, int n, size_t len, int bf16_wire, float prescale,
                   float postscale, float* gsum, float* w, float* m, float* v, float lr,
                   float beta1, float beta2, float eps, long step) {
  if (bf16_wire)
    orc_addn_bf16wire(g, n, len, prescale, gsum);
  else
    orc_addn_f32(g, n, len, prescale, gsum);
  if (postscale != 1.0f) orc_scale_f32(gsum, len, postscale);
  orc_adam_f32(w, m, v, gsum, len, lr, beta1, beta2, eps, step);
}

void orc_step_sgdm(const float* const* g, int n, size_t len, int bf16_wire, float prescale,
                   float postscale, float* gsum, float* w, float* a, float lr, float mu,
                   int nesterov, float wd) {
  if (bf16_wire)
    orc_addn_bf16wire(g, n, len, prescale, gsum);
  else
    orc_addn_f32(g, n, len, prescale, gsum);
  if (postscale != 1.0f) orc_scale_f32(gsum, len, postscale);
  orc_sgdm_f32(w, a, gsum, len, lr, mu, nesterov, wd);
}
