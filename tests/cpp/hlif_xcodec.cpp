// TEST INFRASTRUCTURE: one HLIF (manager) call from a C++ translation unit, linked once against this repo's
// libhipcomp.so and once against the reference's own build (oracle/_ref/libhipcomp_ref.so); the Python test feeds
// the container each one writes to the other (reference container layout:
// src/hipcomp_common_deps/hlif_shared_types.hpp:68-84, src/highlevel/BatchManager.hpp:103-122,213-257).
//
//   hlif_xcodec compress lz4|cascaded <chunk> <raw file> <container file>
//   hlif_xcodec decompress lz4|cascaded <chunk> <container file> <raw file>      (manager of the named codec)
//   hlif_xcodec workspace                                                       (unit test of b200comp::Workspace)
#include "hipcomp/cascaded.hpp"
#include "hipcomp/lz4.hpp"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <memory>
#include <string>
#include <vector>

#ifdef XCODEC_WORKSPACE_TEST
#include "workspace.hpp"
#endif

#define CK(x)                                                                                     \
  do {                                                                                            \
    cudaError_t e_ = (x);                                                                         \
    if (e_ != cudaSuccess) {                                                                      \
      std::fprintf(stderr, "%s -> %s\n", #x, cudaGetErrorString(e_));                             \
      std::exit(2);                                                                               \
    }                                                                                             \
  } while (0)

static std::vector<uint8_t> read_file(const char* p)
{
  std::ifstream f(p, std::ios::binary);
  return std::vector<uint8_t>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

static std::unique_ptr<hipcomp::hipcompManagerBase> make(const std::string& codec, size_t chunk, cudaStream_t st)
{
  if (codec == "lz4") return std::make_unique<hipcomp::LZ4Manager>(chunk, HIPCOMP_TYPE_CHAR, st);
  hipcompBatchedCascadedOpts_t o = hipcompBatchedCascadedDefaultOpts;
  o.type = HIPCOMP_TYPE_INT;
  o.chunk_size = chunk;
  return std::make_unique<hipcomp::CascadedManager>(o, st);
}

#ifdef XCODEC_WORKSPACE_TEST
// reserve / alignment / exhaustion, as the reference tests its TempSpaceBroker (src/test/TempSpaceBroker_test.cpp)
static int workspace_test()
{
  using b200comp::Workspace;
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  int bad = 0;
#define EXPECT(c)                                                         \
  if (!(c)) {                                                             \
    std::fprintf(stderr, "workspace: %s failed (line %d)\n", #c, __LINE__); \
    ++bad;                                                                \
  }
  {
    alignas(64) static uint8_t host[1024];
    Workspace w(host + 1, 1000, 1);
    uint8_t* a = w.reserve<uint8_t>(3);
    EXPECT(a == host + 1);
    double* d = w.reserve<double>(2);
    EXPECT(reinterpret_cast<uintptr_t>(d) % alignof(double) == 0 && reinterpret_cast<uint8_t*>(d) >= a + 3);
    EXPECT(w.spaceLeft() == 1000 - (size_t)(reinterpret_cast<uint8_t*>(d) + 16 - (host + 1)));
    bool threw = false;
    try {
      w.reserve<uint8_t>(w.spaceLeft() + 1);
    } catch (const std::runtime_error&) {
      threw = true;
    }
    EXPECT(threw);
    uint8_t* rest = w.reserve<uint8_t>(w.spaceLeft());
    EXPECT(rest != nullptr && w.spaceLeft() == 0);
    w.rewind();
    EXPECT(w.reserve<uint8_t>(1) == host + 1);
    Workspace w16(host + 1, 1000); // default: at least 16-byte alignment
    EXPECT(reinterpret_cast<uintptr_t>(w16.reserve<uint8_t>(1)) % 16 == 0);
    Workspace empty;
    threw = false;
    try {
      empty.reserve<int>(1);
    } catch (const std::runtime_error&) {
      threw = true;
    }
    EXPECT(threw && !empty.valid());
  }
  {
    // owned scratch comes from the library's private pool: the application's default pool keeps its threshold
    cudaMemPool_t def;
    CK(cudaDeviceGetDefaultMemPool(&def, 0));
    unsigned long long before = 0, after = 0;
    CK(cudaMemPoolGetAttribute(def, cudaMemPoolAttrReleaseThreshold, &before));
    Workspace w(1 << 20, st);
    EXPECT(w.valid() && w.owned() && w.size() == (1u << 20));
    int* p = w.reserve<int>(1000);
    CK(cudaMemsetAsync(p, 0, 4000, st));
    w.release();
    EXPECT(!w.valid());
    w.acquire(2 << 20, st);
    EXPECT(w.valid() && w.size() == (2u << 20));
    alignas(16) static uint8_t other[256];
    w.adopt(other, sizeof other);
    EXPECT(!w.owned() && w.data() == other);
    CK(cudaStreamSynchronize(st));
    CK(cudaMemPoolGetAttribute(def, cudaMemPoolAttrReleaseThreshold, &after));
    EXPECT(before == after);
  }
  std::printf("workspace: %s\n", bad ? "FAILED" : "ok");
  return bad ? 1 : 0;
}
#endif

int main(int argc, char** argv)
{
#ifdef XCODEC_WORKSPACE_TEST
  if (argc >= 2 && std::string(argv[1]) == "workspace") return workspace_test();
#endif
  if (argc < 7) {
    std::fprintf(stderr, "usage: hlif_xcodec compress|decompress lz4|cascaded <chunk> <in> <out> <raw size for decompress or 0>\n");
    return 2;
  }
  const std::string mode = argv[1], codec = argv[2];
  const size_t chunk = std::strtoull(argv[3], nullptr, 10);
  std::vector<uint8_t> in = read_file(argv[4]);
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  auto m = make(codec, chunk, st);
  uint8_t *d_in, *d_out;
  CK(cudaMalloc(&d_in, in.size() + 64));
  CK(cudaMemcpy(d_in, in.data(), in.size(), cudaMemcpyHostToDevice));
  std::vector<uint8_t> out;
  if (mode == "compress") {
    hipcomp::CompressionConfig cfg = m->configure_compression(in.size());
    CK(cudaMalloc(&d_out, cfg.max_compressed_buffer_size));
    CK(cudaMemset(d_out, 0, cfg.max_compressed_buffer_size));
    m->compress(d_in, d_out, cfg);
    CK(cudaStreamSynchronize(st));
    const size_t n = m->get_compressed_output_size(d_out);
    out.resize(n);
    CK(cudaMemcpy(out.data(), d_out, n, cudaMemcpyDeviceToHost));
  } else {
    hipcomp::DecompressionConfig cfg = m->configure_decompression(d_in);
    CK(cudaStreamSynchronize(st));
    CK(cudaMalloc(&d_out, cfg.decomp_data_size + 64));
    m->decompress(d_out, d_in, cfg);
    CK(cudaStreamSynchronize(st));
    if (*cfg.get_status() != hipcompSuccess) {
      std::fprintf(stderr, "decompress status %d\n", (int)*cfg.get_status());
      return 3;
    }
    out.resize(cfg.decomp_data_size);
    CK(cudaMemcpy(out.data(), d_out, out.size(), cudaMemcpyDeviceToHost));
  }
  std::ofstream f(argv[5], std::ios::binary);
  f.write(reinterpret_cast<const char*>(out.data()), (std::streamsize)out.size());
  return 0;
}
