/*
 * BENCH INFRASTRUCTURE (not product code): times the CPU codecs of the reference arm / cpu_baseline of
 * bench.py over the chunks of one buffer, with plain pthreads and clock_gettime, so that the CPU number is
 * not limited by a Python harness.
 *
 *   codec 0: LZ4 block codec of the system liblz4.so.1 (LZ4_compress_default / LZ4_decompress_safe), loaded with
 *            dlopen because the image ships no lz4.h
 *   codec 2: Cascaded, oracle/cascaded_oracle.c (the CPU restatement of the reference algorithm; there is no
 *            other CPU implementation)
 * Snappy has no C-callable library in this image (pyarrow bundles it with hidden symbols); bench.py times it
 * through pyarrow, which releases the GIL inside the codec.
 *
 * Work is dealt dynamically, 16 chunks at a time, to `threads` workers; compress and decompress are timed
 * separately, each over `passes` passes of the whole buffer.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <pthread.h>
#include <stdatomic.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

size_t casc_oracle_compress(const void* in_v, size_t in_bytes, void* out_v, int type, int num_RLEs, int num_deltas,
                            int use_bp);
int casc_oracle_decompress(const void* comp_v, size_t comp_bytes, void* out_v, size_t out_cap, int w, size_t* actual);

typedef int (*lz4_fn)(const char*, char*, int, int);

typedef struct {
  int codec, phase; /* phase 0: compress, 1: decompress */
  const uint8_t* raw;
  uint8_t *comp, *back;
  size_t nchunks, chunk, nbytes, stride;
  size_t* sizes;
  lz4_fn lz4_comp, lz4_dec;
  int casc_type, casc_rle, casc_delta, casc_bp, casc_width;
  atomic_size_t next;
  atomic_int failed;
} job_t;

static void* worker(void* arg)
{
  job_t* j = (job_t*)arg;
  for (;;) {
    const size_t lo = atomic_fetch_add(&j->next, 16);
    if (lo >= j->nchunks) break;
    const size_t hi = lo + 16 < j->nchunks ? lo + 16 : j->nchunks;
    for (size_t i = lo; i < hi; ++i) {
      const size_t off = i * j->chunk;
      const size_t len = off + j->chunk <= j->nbytes ? j->chunk : j->nbytes - off;
      uint8_t* c = j->comp + i * j->stride;
      if (j->phase == 0) {
        if (j->codec == 0) {
          const int r = j->lz4_comp((const char*)j->raw + off, (char*)c, (int)len, (int)j->stride);
          if (r <= 0) atomic_store(&j->failed, 1);
          j->sizes[i] = (size_t)r;
        } else {
          j->sizes[i] = casc_oracle_compress(j->raw + off, len, c, j->casc_type, j->casc_rle, j->casc_delta, j->casc_bp);
        }
      } else {
        if (j->codec == 0) {
          if (j->lz4_dec((const char*)c, (char*)j->back + off, (int)j->sizes[i], (int)len) != (int)len)
            atomic_store(&j->failed, 1);
        } else {
          size_t actual = 0;
          if (casc_oracle_decompress(c, j->sizes[i], j->back + off, len, j->casc_width, &actual) != 0 || actual != len)
            atomic_store(&j->failed, 1);
        }
      }
    }
  }
  return NULL;
}

static double now_s(void)
{
  struct timespec t;
  clock_gettime(CLOCK_MONOTONIC, &t);
  return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec;
}

static double run_phase(job_t* j, int phase, int threads, int passes)
{
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
  j->phase = phase;
  const double t0 = now_s();
  for (int p = 0; p < passes; ++p) {
    atomic_store(&j->next, 0);
    for (int t = 0; t < threads; ++t) pthread_create(&th[t], NULL, worker, j);
    for (int t = 0; t < threads; ++t) pthread_join(th[t], NULL);
  }
  const double dt = now_s() - t0;
  free(th);
  return dt;
}

/*
 * out[0] = seconds of `passes` compress passes, out[1] = seconds of `passes` decompress passes,
 * out[2] = compressed bytes of one pass.  Returns 0, or a negative code (-1 liblz4 missing, -2 out of memory,
 * -3 a codec call failed, -4 round trip differs).
 * casc_opts = {type id, num_RLEs, num_deltas, use_bp, element width in bytes} (ignored for LZ4).
 */
int cpu_bench_run(int codec, const uint8_t* raw, size_t nbytes, size_t chunk, int threads, int passes,
                  const int* casc_opts, double* out)
{
  job_t j;
  memset(&j, 0, sizeof j);
  j.codec = codec;
  j.raw = raw;
  j.nbytes = nbytes;
  j.chunk = chunk;
  j.nchunks = (nbytes + chunk - 1) / chunk;
  if (codec == 0) {
    void* h = dlopen("liblz4.so.1", RTLD_NOW);
    if (!h) return -1;
    j.lz4_comp = (lz4_fn)dlsym(h, "LZ4_compress_default");
    j.lz4_dec = (lz4_fn)dlsym(h, "LZ4_decompress_safe");
    if (!j.lz4_comp || !j.lz4_dec) return -1;
    j.stride = chunk + chunk / 255 + 16; /* LZ4_compressBound */
  } else {
    j.casc_type = casc_opts[0], j.casc_rle = casc_opts[1], j.casc_delta = casc_opts[2], j.casc_bp = casc_opts[3];
    j.casc_width = casc_opts[4];
    j.stride = chunk + 64;
  }
  j.comp = (uint8_t*)malloc(j.nchunks * j.stride);
  j.back = (uint8_t*)malloc(nbytes ? nbytes : 1);
  j.sizes = (size_t*)calloc(j.nchunks ? j.nchunks : 1, sizeof(size_t));
  int rc = 0;
  if (!j.comp || !j.back || !j.sizes) {
    rc = -2;
  } else {
    if (threads < 1) threads = 1;
    out[0] = run_phase(&j, 0, threads, passes);
    out[1] = run_phase(&j, 1, threads, passes);
    size_t total = 0;
    for (size_t i = 0; i < j.nchunks; ++i) total += j.sizes[i];
    out[2] = (double)total;
    if (atomic_load(&j.failed)) rc = -3;
    else if (memcmp(j.back, raw, nbytes) != 0) rc = -4;
  }
  free(j.comp);
  free(j.back);
  free(j.sizes);
  return rc;
}
