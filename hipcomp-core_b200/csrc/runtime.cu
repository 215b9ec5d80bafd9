// b200comp — small runtime utilities shared by the codecs.
#include "common.cuh"

#include <atomic>
#include <cstdlib>
#include <cstring>

namespace b200comp {

int sm_count()
{
  static std::atomic<int> cache[64];
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  int v = cache[dev].load(std::memory_order_relaxed);
  if (v == 0) {
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    cache[dev].store(v, std::memory_order_relaxed);
  }
  return v;
}

int group_lanes_from_env(const char* name)
{
  const char* v = std::getenv(name);
  if (!v) return 0;
  const int g = std::atoi(v);
  return (g == 4 || g == 8 || g == 16 || g == 32) ? g : 0;
}

int decoder_variant_from_env(const char* name)
{
  const char* v = std::getenv(name);
  if (!v || !*v) return kDecoderDefault;
  if (std::strcmp(v, "pipe") == 0) return kDecoderPipeline;
  if (std::strcmp(v, "round") == 0) return kDecoderRound;
  if (std::strcmp(v, "win") == 0) return kDecoderWindow;
  const int g = std::atoi(v);
  return (g == 4 || g == 8 || g == 16 || g == 32) ? g : kDecoderDefault;
}

} // namespace b200comp
