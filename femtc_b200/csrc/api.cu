// C ABI of libfemtc_b200.so (declared in include/femtc_b200.h): argument validation in the
// reference's own order, dispatch between the tiled TMA kernels and the generic kernels, and
// the host-pointer path (chunked, pipelined H2D -> kernel -> D2H).
#include "common.cuh"

#include <nvtx3/nvToolsExt.h> // header-only NVTX v3: no-ops (one pointer test) unless a profiler injected itself

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <thread>
#include <vector>

namespace femtc
{

// ---- error state ---------------------------------------------------------------------------
static thread_local std::string t_last_error;
void set_error(const std::string& msg) { t_last_error = msg; }

int cuda_fail(cudaError_t e, const char* what, const char* file, int line)
{
  char buf[512];
  std::snprintf(buf, sizeof buf, "CUDA error %d (%s) in %s at %s:%d", (int)e, cudaGetErrorString(e),
                what, file, line);
  set_error(buf);
  std::fprintf(stderr, "femtc_b200: %s\n", buf);
  (void)cudaGetLastError(); // clear non-sticky errors
  return FEMTC_ERR_CUDA;
}

// ---- NVTX ranges -------------------------------------------------------------------------------
// The reference wraps its timed loop in a LIKWID marker region named "mass" + typeid(T).name(), i.e. "massd" /
// "massf" (bench_mass.cpp:15-24, 41-47, 94, 108-109; "stiffness.." in bench_stiffness.cpp).  Here every operator
// entry point of the C ABI opens an NVTX range of that same name around its own work (host-pointer entries: the
// whole synchronous call incl. copies; device-pointer entries: the enqueue), so a timeline (nsys / ncu --nvtx)
// shows the regions the LIKWID markers would.  Compiled in unconditionally.
struct NvtxRange
{
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};
template <typename T>
static const char* region(const char* f64_name, const char* f32_name)
{
  return sizeof(T) == 8 ? f64_name : f32_name;
}

static std::atomic<long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static thread_local bool t_dry_run = false;
static thread_local LaunchRecord t_record;
bool dry_run() { return t_dry_run; }
LaunchRecord& launch_record() { return t_record; }

// ---- options ---------------------------------------------------------------------------------
static std::atomic<long long> g_opt_kernel{0};
static std::atomic<long long> g_opt_host_chunk_mb{64};
static long long env_int(const char* name, long long lo, long long hi, long long dflt);
static std::atomic<long long> g_opt_variant{0};
static long long env_flag(const char* name)
{
  const char* v = std::getenv(name);
  return v && *v && *v != '0' ? 1 : 0;
}
static std::atomic<long long> g_opt_host_register{env_flag("FEMTC_HOST_REGISTER")};
static long long env_int(const char* name, long long lo, long long hi, long long dflt)
{
  const char* v = std::getenv(name);
  if (!v || !*v)
    return dflt;
  const long long x = std::atoll(v);
  return x < lo || x > hi ? dflt : x;
}
static std::atomic<long long> g_opt_host_devices{env_int("FEMTC_HOST_DEVICES", 1, 64, 1)};
static std::atomic<long long> g_opt_host_oversubscribe{env_int("FEMTC_HOST_DEVICES_OVERSUBSCRIBE", 0, 1, 0)};
static std::atomic<long long> g_opt_tc{env_int("FEMTC_TC", -1, 1, -1)}; // env: initial value of option "tc" (A/B runs of whole programs)

long long opt_variant() { return g_opt_variant.load(); }
long long opt_tc() { return g_opt_tc.load(); }

// ---- device bookkeeping ---------------------------------------------------------------------
static int g_sm_count[64];
static int g_usable_devices = -1;
static std::once_flag g_probe_once;

static void probe_devices()
{
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess)
  {
    (void)cudaGetLastError();
    n = 0;
  }
  int usable = 0;
  for (int d = 0; d < n && d < 64; ++d)
  {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, d) != cudaSuccess)
      break;
    if (p.major != 10) // the library holds sm_100a code only
      break;
    g_sm_count[d] = p.multiProcessorCount;
    ++usable;
  }
  g_usable_devices = usable;
}

static int require_device()
{
  std::call_once(g_probe_once, probe_devices);
  if (g_usable_devices <= 0)
  {
    set_error("no usable sm_100 (B200) device: libfemtc_b200 has no CPU fallback");
    return FEMTC_ERR_CUDA;
  }
  return FEMTC_OK;
}

int sm_count()
{
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= 64 || g_sm_count[d] == 0)
    return 148;
  return g_sm_count[d];
}

// ---- validation, in the order the reference dispatches (Nb, Mb, order, degree) ----------------
static int check_knobs(int degree, int order, int Mb, int Nb)
{
  auto ok_block = [](int b) { return b == 0 || b == 4 || b == 8 || b == 16; };
  if (!ok_block(Nb) || !ok_block(Mb))
  {
    set_error("Block size not supported");
    return FEMTC_ERR_BLOCK;
  }
  if (order < 0 || order > 5)
  {
    set_error("order not supported");
    return FEMTC_ERR_ORDER;
  }
  if (degree < 1 || degree > 25)
  {
    set_error("degree not supported");
    return FEMTC_ERR_DEGREE;
  }
  return FEMTC_OK;
}

// ---- device-pointer operators --------------------------------------------------------------------
template <typename T>
static int finish_checksum(const T* dW, size_t num_cells, size_t nd, size_t done, double* partials, int count,
                           double* d_out4, cudaStream_t stream)
{
  // `count` triples cover the first `done` cells (0: the kernel family has no fused epilogue); the rest of W -- the
  // ragged tail, or everything -- goes through the stand-alone first stage, then one second stage sums all triples
  const size_t covered = count > 0 ? done : 0;
  int rc = FEMTC_OK;
  if (covered < num_cells)
  {
    if ((rc = launch_checksum_stage1<T>(dW + covered * nd, (num_cells - covered) * nd, partials + 3 * count, stream)))
      return rc;
    count += kChecksumStage1Triples;
  }
  return launch_checksum_stage2(partials, count, num_cells * nd, d_out4, stream);
}

template <typename T>
static int mass_device(const T* phi, size_t phi_len, const T* dU, T* dW, const T* ddetJ,
                       size_t num_cells, int degree, cudaStream_t stream, double* d_checksum4 = nullptr)
{
  NvtxRange nvtx(region<T>("massd:enqueue", "massf:enqueue"));
  if (degree < 1 || degree > 25)
  {
    set_error("degree not supported");
    return FEMTC_ERR_DEGREE;
  }
  const size_t n2 = (size_t)(degree + 1) * (degree + 1);
  if (!phi || (num_cells && (!dU || !dW || !ddetJ)))
  {
    set_error("null pointer argument");
    return FEMTC_ERR_ARG;
  }
  if (phi_len < n2)
  {
    set_error("phi shorter than (degree+1)^2");
    return FEMTC_ERR_LENGTH;
  }
  int rc = require_device();
  if (rc)
    return rc;
  MassArgs<T> a{phi, dU, dW, ddetJ, num_cells, degree, stream};
  double* partials = nullptr;
  int cs_count = 0;
  if (d_checksum4 && !dry_run())
  {
    if ((rc = checksum_scratch(&partials)))
      return rc;
    a.cs_partials = partials, a.cs_capacity = kChecksumFused, a.cs_count = &cs_count;
  }
  size_t done = 0;
  const long long force = g_opt_kernel.load();
  if (force != 1 && degree <= 8)
  {
    long long n = launch_mass_cell<T>(a);
    if (n == 0)
      n = launch_mass_tiled<T>(a);
    if (n == 0)
      n = launch_mass_tc<T>(a);
    if (n == 0)
      n = launch_mass_column<T>(a);
    if (n < 0)
      return (int)-n;
    done = (size_t)n;
    if (force == 2 && done == 0 && num_cells)
    {
      set_error("tiled kernel forced but not applicable (alignment or size)");
      return FEMTC_ERR_ARG;
    }
  }
  if (done < num_cells && !(dry_run() && done)) // (a dry run reports the kernel of the leading cells, not the tail's)
  {
    const size_t nd = n2 * (degree + 1);
    MassArgs<T> t{phi, dU + done * nd, dW + done * nd, ddetJ + done * nd, num_cells - done, degree, stream};
    rc = launch_mass_generic<T>(t);
  }
  if (rc == FEMTC_OK && partials)
    rc = finish_checksum<T>(dW, num_cells, n2 * (degree + 1), done, partials, cs_count, d_checksum4, stream);
  return rc;
}

static std::once_flag g_short_phi_once;

template <typename T>
static int stiffness_device(const T* phi, size_t phi_len, const T* dU, T* dW, const T* dG,
                            size_t num_cells, int degree, int mode, cudaStream_t stream, double* d_checksum4 = nullptr)
{
  NvtxRange nvtx(region<T>("stiffnessd:enqueue", "stiffnessf:enqueue"));
  if (degree < 1 || degree > 25)
  {
    set_error("degree not supported");
    return FEMTC_ERR_DEGREE;
  }
  if (mode != FEMTC_STIFFNESS_REF_EXACT && mode != FEMTC_STIFFNESS_FULL)
  {
    set_error("unknown stiffness mode");
    return FEMTC_ERR_MODE;
  }
  const size_t n2 = (size_t)(degree + 1) * (degree + 1);
  if (!phi || (num_cells && (!dU || !dW || !dG)))
  {
    set_error("null pointer argument");
    return FEMTC_ERR_ARG;
  }
  if (phi_len < n2)
  {
    set_error("phi shorter than (degree+1)^2");
    return FEMTC_ERR_LENGTH;
  }
  const T* A = phi;
  const T* D = phi + n2;
  if (phi_len < 2 * n2)
  {
    // bench_stiffness.cpp:74 allocates (p+1)^2 values but stiffness.hpp:45 reads the derivative
    // matrix from phi[(p+1)^2 ...): never read past what the caller gave us.
    std::call_once(g_short_phi_once, [] {
      std::fprintf(stderr, "femtc_b200: stiffness called with phi of (P+1)^2 values; the operator needs "
                           "2(P+1)^2 = [basis | derivative]. Using derivative := basis (no out-of-bounds read).\n");
    });
    D = A;
  }
  int rc = require_device();
  if (rc)
    return rc;
  StiffnessArgs<T> a{A, D, dU, dW, dG, num_cells, degree, mode, stream};
  double* partials = nullptr;
  int cs_count = 0;
  if (d_checksum4 && !dry_run())
  {
    if ((rc = checksum_scratch(&partials)))
      return rc;
    a.cs_partials = partials, a.cs_capacity = kChecksumFused, a.cs_count = &cs_count;
  }
  size_t done = 0;
  const long long force = g_opt_kernel.load();
  if (force != 1 && degree <= 8)
  {
    long long n = launch_stiffness_cell<T>(a);
    if (n == 0)
      n = launch_stiffness_tiled<T>(a);
    if (n == 0)
      n = launch_stiffness_tc<T>(a);
    if (n == 0)
      n = launch_stiffness_column<T>(a);
    if (n < 0)
      return (int)-n;
    done = (size_t)n;
    if (force == 2 && done == 0 && num_cells)
    {
      set_error("tiled kernel forced but not applicable (alignment or size)");
      return FEMTC_ERR_ARG;
    }
  }
  if (done < num_cells && !(dry_run() && done))
  {
    const size_t nd = n2 * (degree + 1);
    StiffnessArgs<T> t{A, D, dU + done * nd, dW + done * nd, dG + done * nd * 6, num_cells - done, degree, mode, stream};
    rc = launch_stiffness_generic<T>(t);
  }
  if (rc == FEMTC_OK && partials)
    rc = finish_checksum<T>(dW, num_cells, n2 * (degree + 1), done, partials, cs_count, d_checksum4, stream);
  return rc;
}

// ---- host-pointer path: chunked three-deep pipeline ------------------------------------------------
// Each slot owns a stream and device buffers for one chunk of cells.  Per chunk:
//   H2D(U), H2D(detJ|G) -> operator kernel -> D2H(W), all on the slot's stream, so chunk i+1's
// uploads and chunk i-1's download overlap chunk i's kernel (separate copy engines per direction).
// Page-locked caller buffers (femtc_host_alloc / cudaHostAlloc / cudaHostRegister) are DMA'd in
// place; pageable buffers go through the driver's staging path.
struct Slot
{
  cudaStream_t stream = nullptr;
  void* dU = nullptr;
  void* dG = nullptr;
  void* dW = nullptr;
  size_t capU = 0, capG = 0, capW = 0;
};

struct DeviceCache
{
  std::mutex mu;
  Slot slots[3];
  bool streams_ready = false;
  void* scratch = nullptr; // grow-only staging for the small GEMM / contraction host entries
  size_t scratch_cap = 0;
};
static DeviceCache g_cache[64];

static int ensure(void*& p, size_t& cap, size_t bytes)
{
  if (cap >= bytes)
    return FEMTC_OK;
  if (p)
    FEMTC_CUDA(cudaFree(p));
  p = nullptr;
  cap = 0;
  FEMTC_CUDA(cudaMalloc(&p, bytes));
  cap = bytes;
  return FEMTC_OK;
}

// ---- opt-in page-locking of the caller's (pageable) buffers ---------------------------------------------
// The reference's drivers hold their fields in std::vector and call the operator nreps times on the same
// buffers (bench_mass.cpp:95-105).  Pageable memory is staged by the driver at a fraction of the PCIe rate;
// with option "host_register" = 1 (or FEMTC_HOST_REGISTER=1 in the environment) a buffer of >= 1 MB is
// page-locked in place the first time it is seen (cudaHostRegister, portable: every device may DMA it) and the
// registration is remembered by address, so later calls DMA straight from / into the caller's memory.  OFF by
// default: a registration must be dropped before the memory is freed (femtc_release_cache() does that), which a
// drop-in library cannot know about.
static std::mutex g_reg_mu;
static std::map<uintptr_t, size_t> g_registered; // base -> bytes
// host-pointer calls between their registration step and their return (guarded by g_reg_mu): while any is in flight no
// registration is dropped, because its copies may be running through it (another thread's call on an overlapping slice)
static int g_host_calls_in_flight = 0;

static bool page_locked(const void* p)
{
  cudaPointerAttributes attr{};
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess)
  {
    (void)cudaGetLastError();
    return false;
  }
  return attr.type != cudaMemoryTypeUnregistered;
}

// caller holds g_reg_mu
static void maybe_register(const void* ptr, size_t bytes)
{
  if (!g_opt_host_register.load() || !ptr || bytes < (size_t(1) << 20))
    return;
  const uintptr_t base = reinterpret_cast<uintptr_t>(ptr), end = base + bytes;
  static const bool debug = env_flag("FEMTC_DEBUG_REGISTER"); // development: log every decision on stderr
  // One of our own registrations that covers the range: done.  One that only overlaps it is stale (the caller
  // freed the buffer and got the address again with another size) or too short: drop it and register afresh.
  for (auto it = g_registered.begin(); it != g_registered.end();)
  {
    const uintptr_t b = it->first, e = b + it->second;
    if (b <= base && end <= e)
    {
      if (debug)
        std::fprintf(stderr, "femtc_b200: %p +%zu covered by an earlier registration\n", ptr, bytes);
      return;
    }
    if (b < end && base < e)
    {
      if (g_host_calls_in_flight > 0)
      {
        // another call may be copying through that entry right now: leave it, serve this range pageable this time
        if (debug)
          std::fprintf(stderr, "femtc_b200: %p +%zu overlaps own entry %p +%zu in use: left alone\n", ptr, bytes, (void*)b, it->second);
        return;
      }
      if (debug)
        std::fprintf(stderr, "femtc_b200: %p +%zu overlaps own entry %p +%zu: dropping it\n", ptr, bytes, (void*)b, it->second);
      if (cudaHostUnregister(reinterpret_cast<void*>(b)) != cudaSuccess)
        (void)cudaGetLastError();
      it = g_registered.erase(it);
    }
    else
      ++it;
  }
  // page-locked by somebody else (cudaHostAlloc, a foreign cudaHostRegister, device or managed memory): the whole
  // range has to be, first byte to last; a partly covered range cannot be completed in place and stays as it is
  const bool first = page_locked(ptr), last = page_locked(reinterpret_cast<const char*>(ptr) + bytes - 1);
  if (first || last)
  {
    if (debug)
      std::fprintf(stderr, "femtc_b200: register %p +%zu: already page-locked (first %d, last %d)\n", ptr, bytes, first, last);
    return;
  }
  const cudaError_t e = cudaHostRegister(const_cast<void*>(ptr), bytes, cudaHostRegisterPortable);
  if (debug)
    std::fprintf(stderr, "femtc_b200: register %p +%zu: %s\n", ptr, bytes, cudaGetErrorString(e));
  if (e == cudaSuccess)
    g_registered[base] = bytes;
  else
    (void)cudaGetLastError(); // overlaps a foreign registration, or not registrable: stay on the pageable path
}

static void unregister_all()
{
  std::lock_guard<std::mutex> lock(g_reg_mu);
  for (auto& r : g_registered)
    if (cudaHostUnregister(reinterpret_cast<void*>(r.first)) != cudaSuccess)
      (void)cudaGetLastError();
  g_registered.clear();
}

// Chunk and share boundaries of the host-pointer path fall on multiples of cut_cells(n1) cells: 16-byte aligned in every
// array for any dtype (TMA path), and a whole number of tiles of every kernel family that serves that n1 (slab rounds
// of 32/n1 cells and pairs of them, 32-cell rounds of the one-thread-per-cell kernels, column tiles), so each cell is
// served by the same kernel as in an unchunked call and the result depends neither on the chunk size nor on the fan-out.
static size_t cut_cells(size_t nd)
{
  switch (nd)
  {
  case 8: return 64;     // n1 = 2: rounds of 16 / 32, column tile 64
  case 27: return 1120;  // n1 = 3: rounds of 10 / 32, column tile 28
  case 125: return 240;  // n1 = 5: pairs of 6-cell rounds, column tiles 8 / 10
  case 216: return 560;  // n1 = 6: rounds of 5, column tiles 7 / 8 / 14 / 16
  default: return 16;    // n1 = 4, 7, 8, 9: tiles of 2 / 4 / 8; generic kernel: any
  }
}

// One device's share of a host-pointer call: the calling thread's current device sweeps `num_cells` cells in chunks.
// `geom_dev` != nullptr: the geometry already lives in HBM (operator handle, below); only U and W travel.
// Every exit, error or not, goes through the epilogue that drains the three slot streams: when the call returns
// nothing is in flight any more, in particular no D2H copy into the caller's W.
template <typename T, typename Launch>
static int host_pipeline_device(const T* U, T* W, const T* geom, int geom_width, size_t num_cells, size_t nd,
                                Launch&& launch, const T* geom_dev)
{
  int dev = 0;
  FEMTC_CUDA(cudaGetDevice(&dev));
  DeviceCache& cache = g_cache[dev];
  std::lock_guard<std::mutex> lock(cache.mu);
  if (!cache.streams_ready)
  {
    for (auto& s : cache.slots)
      FEMTC_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    cache.streams_ready = true;
  }
  const size_t bytes_per_cell = nd * sizeof(T) * (1 + (geom_dev ? 0 : geom_width));
  size_t chunk_cells = std::max<size_t>(1, ((size_t)g_opt_host_chunk_mb.load() << 20) / bytes_per_cell);
  // chunk boundaries on whole tiles of every kernel family (cut_cells above)
  const size_t kCutCells = cut_cells(nd);
  chunk_cells = std::max<size_t>(kCutCells, chunk_cells / kCutCells * kCutCells);
  chunk_cells = std::min(chunk_cells, std::max<size_t>(num_cells, 1));

  int rc = FEMTC_OK;
  auto cuda_ok = [&rc](cudaError_t e, const char* what, int line)
  {
    if (e != cudaSuccess && rc == FEMTC_OK)
      rc = cuda_fail(e, what, __FILE__, line);
    return e == cudaSuccess;
  };
  size_t i = 0;
  for (size_t c0 = 0; c0 < num_cells && rc == FEMTC_OK; c0 += chunk_cells, ++i)
  {
    const size_t n = std::min(chunk_cells, num_cells - c0);
    Slot& s = cache.slots[i % 3];
    // the slot's previous chunk (three chunks ago) must have drained before its buffers are reused
    if (!cuda_ok(cudaStreamSynchronize(s.stream), "cudaStreamSynchronize(slot)", __LINE__))
      break;
    if ((rc = ensure(s.dU, s.capU, chunk_cells * nd * sizeof(T))))
      break;
    if (!geom_dev && (rc = ensure(s.dG, s.capG, chunk_cells * nd * geom_width * sizeof(T))))
      break;
    if ((rc = ensure(s.dW, s.capW, chunk_cells * nd * sizeof(T))))
      break;
    if (!cuda_ok(cudaMemcpyAsync(s.dU, U + c0 * nd, n * nd * sizeof(T), cudaMemcpyHostToDevice, s.stream), "H2D(U)", __LINE__))
      break;
    if (!geom_dev && !cuda_ok(cudaMemcpyAsync(s.dG, geom + c0 * nd * geom_width, n * nd * geom_width * sizeof(T),
                                              cudaMemcpyHostToDevice, s.stream), "H2D(geometry)", __LINE__))
      break;
    rc = launch((const T*)s.dU, (T*)s.dW, geom_dev ? geom_dev + c0 * nd * geom_width : (const T*)s.dG, n, s.stream);
    if (rc)
      break;
    if (!cuda_ok(cudaMemcpyAsync(W + c0 * nd, s.dW, n * nd * sizeof(T), cudaMemcpyDeviceToHost, s.stream), "D2H(W)", __LINE__))
      break;
  }
  for (auto& s : cache.slots)
    cuda_ok(cudaStreamSynchronize(s.stream), "cudaStreamSynchronize(slot)", __LINE__);
  return rc;
}

// The host-pointer call.  By default the calling thread's current device takes the whole cell range.  With option
// "host_devices" = n > 1 (env FEMTC_HOST_DEVICES) the range is cut into n contiguous shares (16-cell boundaries),
// one per device 0..n-1, each swept by its own thread through that device's three-slot pipeline: the action is
// cell-local, so the shares are independent and the result is bit-identical to the one-device call, while the
// copies use n PCIe links instead of one.  Not used with a resident-geometry handle (it lives on one device).
template <typename T, typename Launch>
static int host_pipeline(const T* U, T* W, const T* geom, int geom_width, size_t num_cells, size_t nd,
                         Launch&& launch, const T* geom_dev = nullptr)
{
  struct InFlight // registration step and in-flight count under one lock; the count is given back on every exit
  {
    InFlight(const T* U, T* W, const T* geom, size_t bU, size_t bG)
    {
      std::lock_guard<std::mutex> lock(g_reg_mu);
      maybe_register(U, bU);
      maybe_register(W, bU);
      if (geom)
        maybe_register(geom, bG);
      ++g_host_calls_in_flight;
    }
    ~InFlight()
    {
      std::lock_guard<std::mutex> lock(g_reg_mu);
      --g_host_calls_in_flight;
    }
  } in_flight(U, W, geom_dev ? nullptr : geom, num_cells * nd * sizeof(T), num_cells * nd * geom_width * sizeof(T));
  const size_t kCutCells = cut_cells(nd);
  // option "host_devices_oversubscribe" / FEMTC_HOST_DEVICES_OVERSUBSCRIBE=1 (testing on a box with fewer GPUs than shares): the share count is not clamped
  // and share r runs on device r % visible; shares of one device take turns on its pipeline
  const bool oversubscribe = g_opt_host_oversubscribe.load() != 0;
  const int visible = g_usable_devices;
  int ndev = (int)(oversubscribe ? g_opt_host_devices.load() : std::min<long long>(g_opt_host_devices.load(), visible));
  if (geom_dev || ndev < 2 || visible < 1 || num_cells < (size_t)ndev * kCutCells)
    return host_pipeline_device<T>(U, W, geom, geom_width, num_cells, nd, launch, geom_dev);

  std::vector<int> rcs(ndev, FEMTC_OK);
  std::vector<std::string> errs(ndev);
  std::vector<std::thread> workers;
  auto cut = [&](int r) { return r == ndev ? num_cells : (num_cells * r / ndev) / kCutCells * kCutCells; };
  bool spawn_failed = false;
  for (int r = 0; r < ndev && !spawn_failed; ++r)
  {
    const size_t c0 = cut(r), c1 = cut(r + 1);
    try
    {
      workers.emplace_back([&, r, c0, c1]
    {
      cudaError_t e = cudaSetDevice(r % visible);
      if (e != cudaSuccess)
      {
        rcs[r] = cuda_fail(e, "cudaSetDevice(host_devices share)", __FILE__, __LINE__);
      }
      else if (c1 > c0)
        rcs[r] = host_pipeline_device<T>(U + c0 * nd, W + c0 * nd, geom + c0 * nd * geom_width, geom_width, c1 - c0,
                                         nd, launch, (const T*)nullptr);
      if (rcs[r])
        errs[r] = t_last_error; // the error text is thread-local: hand it to the caller's thread
    });
    }
    catch (const std::exception&) // thread creation failed: the workers already started are joined below
    {
      spawn_failed = true;
    }
  }
  for (auto& w : workers)
    w.join();
  if (spawn_failed)
  {
    set_error("host_devices: could not start a worker thread");
    return FEMTC_ERR_CUDA;
  }
  for (int r = 0; r < ndev; ++r)
    if (rcs[r])
    {
      set_error(errs[r]);
      return rcs[r];
    }
  return FEMTC_OK;
}

// ---- opt-in geometry cache behind the drop-in host-pointer call ----------------------------------
// The unchanged drivers apply the operator nreps times to the SAME detJ / G (bench_mass.cpp:95-105,
// bench_stiffness.cpp:95-105), and the geometry is 1 of 3 (mass) or 6 of 8 (stiffness) words per DoF that cross PCIe on
// every call.  With option "geom_cache" = 1 (FEMTC_GEOM_CACHE=1) the host-pointer entry points keep the geometry array
// in HBM, keyed by (device, host address, bytes) and guarded by a 64-bit hash of 4 KB sampled at fixed strides: a later
// call with the same key and hash moves only U in and W out (the ResidentOperator path, same kernels, same chunk cuts:
// bit-identical).  A changed hash re-uploads; femtc_release_cache() drops everything.  OFF by default: a drop-in
// library cannot know that the caller rewrote geometry between calls in a place the samples miss -- a caller who
// does that must call femtc_release_cache() (or leave the option off).
static std::atomic<long long> g_opt_geom_cache{env_int("FEMTC_GEOM_CACHE", 0, 1, 0)};
static std::atomic<long long> g_geom_uploads{0};
struct GeomEntry
{
  int device;
  const void* host;
  size_t bytes;
  uint64_t hash;
  void* dev;
};
static std::mutex g_geom_mu;
static std::vector<GeomEntry> g_geom;

static uint64_t sampled_hash(const void* p, size_t bytes)
{
  // 64 blocks of 64 bytes, first block at 0, last block ending at the last whole word; 8-byte words through splitmix64
  const unsigned char* b = static_cast<const unsigned char*>(p);
  const size_t words = bytes / 8;
  uint64_t h = splitmix64(bytes);
  if (words == 0)
    return h;
  const size_t blocks = 64, per = 8;
  for (size_t i = 0; i < blocks; ++i)
  {
    const size_t w0 = words <= per ? 0 : (size_t)((long double)(words - per) * i / (blocks - 1));
    for (size_t k = 0; k < per && w0 + k < words; ++k)
    {
      uint64_t v;
      std::memcpy(&v, b + (w0 + k) * 8, 8);
      h = splitmix64(h ^ v);
    }
  }
  return h;
}

// Device copy of `geom` on the current device, or nullptr when the cache is off / cannot hold it (the caller then takes
// the strict path; never an error).
static const void* cached_geometry(const void* geom, size_t bytes)
{
  if (!g_opt_geom_cache.load() || !geom || bytes == 0)
    return nullptr;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess)
    return nullptr;
  const uint64_t h = sampled_hash(geom, bytes);
  std::lock_guard<std::mutex> lock(g_geom_mu);
  GeomEntry* e = nullptr;
  for (auto& g : g_geom)
    if (g.device == dev && g.host == geom && g.bytes == bytes)
      e = &g;
  if (e && e->hash == h)
    return e->dev;
  if (!e)
  {
    void* d = nullptr;
    if (cudaMalloc(&d, bytes) != cudaSuccess)
    {
      (void)cudaGetLastError(); // no room in HBM for a resident copy: strict path
      return nullptr;
    }
    g_geom.push_back({dev, geom, bytes, 0, d});
    e = &g_geom.back();
  }
  if (cudaMemcpy(e->dev, geom, bytes, cudaMemcpyHostToDevice) != cudaSuccess)
  {
    (void)cudaGetLastError();
    cudaFree(e->dev);
    g_geom.erase(g_geom.begin() + (e - g_geom.data()));
    return nullptr;
  }
  e->hash = h;
  g_geom_uploads.fetch_add(1);
  return e->dev;
}

static void drop_cached_geometry()
{
  std::lock_guard<std::mutex> lock(g_geom_mu);
  int cur = 0;
  const bool have = cudaGetDevice(&cur) == cudaSuccess;
  for (auto& g : g_geom)
  {
    cudaSetDevice(g.device);
    cudaFree(g.dev);
  }
  g_geom.clear();
  if (have)
    cudaSetDevice(cur);
}

// one device serves the call and the geometry may stay resident there
static bool geometry_cacheable()
{
  const long long shares = g_opt_host_oversubscribe.load() ? g_opt_host_devices.load()
                                                           : std::min<long long>(g_opt_host_devices.load(), g_usable_devices);
  return g_opt_geom_cache.load() && shares < 2;
}

template <typename T>
static int mass_host(const T* phi, size_t phi_len, const T* U, T* W, const T* detJ, size_t num_cells,
                     int degree, int order, int Mb, int Nb)
{
  NvtxRange nvtx(region<T>("massd", "massf")); // the LIKWID region name of bench_mass.cpp:41-47
  int rc = check_knobs(degree, order, Mb, Nb);
  if (rc)
    return rc;
  const size_t n1 = degree + 1;
  if (!phi || (num_cells && (!U || !W || !detJ)))
  {
    set_error("null pointer argument");
    return FEMTC_ERR_ARG;
  }
  if (phi_len < n1 * n1)
  {
    set_error("phi shorter than (degree+1)^2");
    return FEMTC_ERR_LENGTH;
  }
  if ((rc = require_device()))
    return rc;
  if (num_cells == 0)
    return FEMTC_OK;
  const T* resident = geometry_cacheable()
                          ? static_cast<const T*>(cached_geometry(detJ, num_cells * n1 * n1 * n1 * sizeof(T))) : nullptr;
  return host_pipeline<T>(U, W, resident ? (const T*)nullptr : detJ, 1, num_cells, n1 * n1 * n1,
                          [&](const T* dU, T* dW, const T* dJ, size_t n, cudaStream_t s)
                          { return mass_device<T>(phi, phi_len, dU, dW, dJ, n, degree, s); },
                          resident);
}

template <typename T>
static int stiffness_host(const T* phi, size_t phi_len, const T* U, T* W, const T* G, size_t num_cells,
                          int degree, int order, int Mb, int Nb, int mode)
{
  NvtxRange nvtx(region<T>("stiffnessd", "stiffnessf")); // bench_stiffness.cpp:41-47
  int rc = check_knobs(degree, order, Mb, Nb);
  if (rc)
    return rc;
  if (mode != FEMTC_STIFFNESS_REF_EXACT && mode != FEMTC_STIFFNESS_FULL)
  {
    set_error("unknown stiffness mode");
    return FEMTC_ERR_MODE;
  }
  const size_t n1 = degree + 1;
  if (!phi || (num_cells && (!U || !W || !G)))
  {
    set_error("null pointer argument");
    return FEMTC_ERR_ARG;
  }
  if (phi_len < n1 * n1)
  {
    set_error("phi shorter than (degree+1)^2");
    return FEMTC_ERR_LENGTH;
  }
  if ((rc = require_device()))
    return rc;
  if (num_cells == 0)
    return FEMTC_OK;
  const T* resident = geometry_cacheable()
                          ? static_cast<const T*>(cached_geometry(G, num_cells * n1 * n1 * n1 * 6 * sizeof(T))) : nullptr;
  return host_pipeline<T>(U, W, resident ? (const T*)nullptr : G, 6, num_cells, n1 * n1 * n1,
                          [&](const T* dU, T* dW, const T* dG, size_t n, cudaStream_t s)
                          { return stiffness_device<T>(phi, phi_len, dU, dW, dG, n, degree, mode, s); },
                          resident);
}

// ---- small host-pointer helpers for the GEMM / contraction entry points ----------------------------
// The reference's tests call these tens of times with tiny operands (tests/test_gemm.cpp:41,52): their device
// staging comes out of one grow-only scratch buffer per device, held under the device's mutex for the call.
struct Scratch
{
  std::unique_lock<std::mutex> lock;
  char* base = nullptr;
  size_t used = 0;
  // sizes[i] bytes each, 256-byte aligned pieces of the device's scratch buffer
  int open(std::initializer_list<size_t> sizes, void** out)
  {
    int dev = 0;
    FEMTC_CUDA(cudaGetDevice(&dev));
    DeviceCache& cache = g_cache[dev];
    lock = std::unique_lock<std::mutex>(cache.mu);
    size_t total = 0;
    for (size_t b : sizes)
      total += (std::max<size_t>(b, 16) + 255) & ~size_t(255);
    int rc = ensure(cache.scratch, cache.scratch_cap, std::max<size_t>(total, size_t(1) << 20));
    if (rc)
      return rc;
    base = static_cast<char*>(cache.scratch);
    for (size_t b : sizes)
    {
      *out++ = base + used;
      used += (std::max<size_t>(b, 16) + 255) & ~size_t(255);
    }
    return FEMTC_OK;
  }
};

template <typename T>
static int gemm_blocked_host(int k, int m, int n, int order, int MB, int NB, const T* a, const T* b, T* c)
{
  if (k < 1 || m < 1 || n < 1 || MB < 0 || NB < 0 || !a || !b || !c)
  {
    set_error("bad gemm_blocked shape or null pointer");
    return FEMTC_ERR_ARG;
  }
  if (order < 0 || order > 5)
  {
    set_error("order not supported");
    return FEMTC_ERR_ORDER;
  }
  int rc = require_device();
  if (rc)
    return rc;
  Scratch sc;
  void* d[3];
  if ((rc = sc.open({sizeof(T) * m * k, sizeof(T) * k * n, sizeof(T) * m * n}, d)))
    return rc;
  FEMTC_CUDA(cudaMemcpy(d[0], a, sizeof(T) * m * k, cudaMemcpyHostToDevice));
  FEMTC_CUDA(cudaMemcpy(d[1], b, sizeof(T) * k * n, cudaMemcpyHostToDevice));
  if ((rc = launch_gemm_blocked<T>(k, m, n, (const T*)d[0], (const T*)d[1], (T*)d[2], nullptr)))
    return rc;
  FEMTC_CUDA(cudaMemcpy(c, d[2], sizeof(T) * m * n, cudaMemcpyDeviceToHost));
  return FEMTC_OK;
}

template <typename T>
static int gemm_strided_host(int k, int m, int nc, int sai, int sap, int sbp, int sbj, int ldc, const T* a,
                             const T* b, T* c)
{
  if (k < 1 || m < 1 || nc < 1 || sai < 1 || sap < 1 || sbp < 1 || sbj < 1 || ldc < nc || !a || !b || !c)
  {
    set_error("bad strided gemm shape or null pointer");
    return FEMTC_ERR_ARG;
  }
  int rc = require_device();
  if (rc)
    return rc;
  // extents actually touched
  const size_t na = (size_t)(m - 1) * sai + (size_t)(k - 1) * sap + 1;
  const size_t nb = (size_t)(k - 1) * sbp + (size_t)(nc - 1) * sbj + 1;
  const size_t ncc = (size_t)(m - 1) * ldc + nc;
  Scratch sc;
  void* d[3];
  if ((rc = sc.open({sizeof(T) * na, sizeof(T) * nb, sizeof(T) * ncc}, d)))
    return rc;
  FEMTC_CUDA(cudaMemcpy(d[0], a, sizeof(T) * na, cudaMemcpyHostToDevice));
  FEMTC_CUDA(cudaMemcpy(d[1], b, sizeof(T) * nb, cudaMemcpyHostToDevice));
  FEMTC_CUDA(cudaMemcpy(d[2], c, sizeof(T) * ncc, cudaMemcpyHostToDevice));
  if ((rc = launch_gemm_strided<T>(k, m, nc, sai, sap, sbp, sbj, ldc, (const T*)d[0], (const T*)d[1], (T*)d[2],
                                   nullptr)))
    return rc;
  // copy back row by row so that the gaps between rows of c (ldc > nc) keep the caller's values
  FEMTC_CUDA(cudaMemcpy2D(c, sizeof(T) * ldc, d[2], sizeof(T) * ldc, sizeof(T) * nc, m, cudaMemcpyDeviceToHost));
  return FEMTC_OK;
}

template <typename T>
static int gemm_micro_host(int k, int m, int nc, int lda, int ldb, int ldc, int order, const T* a,
                           const T* b, T* c)
{
  if (k < 1 || m < 1 || nc < 1 || lda < m || ldb < nc || ldc < nc)
  {
    set_error("bad gemm_micro shape");
    return FEMTC_ERR_ARG;
  }
  if (order < 0 || order > 5)
  {
    set_error("order not supported");
    return FEMTC_ERR_ORDER;
  }
  return gemm_strided_host<T>(k, m, nc, 1, lda, ldb, 1, ldc, a, b, c);
}

// tmm_kernels.cu
template <typename T>
long long launch_tmm_tiled(int Na, int Nb, int Tr, const T* I0, const T* I1, const T* I2, const T* dB, T* dA,
                           size_t batch, cudaStream_t s);

// B, A device pointers; I0..I2 host pointers.  The tiled kernel serves whole tiles of the covered shapes, the
// generic kernel everything else (ragged tail, other shapes, unaligned pointers) on the same stream.
template <typename T>
static int apply_contractions_device(int Na, int Nb, int Tr, const T* I0, const T* I1, const T* I2, const T* dB,
                                     T* dA, size_t batch, cudaStream_t stream)
{
  NvtxRange nvtx(region<T>("apply_contractionsd:enqueue", "apply_contractionsf:enqueue"));
  if (Na < 1 || Nb < 1 || Na > 64 || Nb > 64 || !I0 || !I1 || !I2 || (batch && (!dA || !dB)))
  {
    set_error("bad apply_contractions shape or null pointer");
    return FEMTC_ERR_ARG;
  }
  int rc = require_device();
  if (rc || batch == 0)
    return rc;
  size_t done = 0;
  if (g_opt_kernel.load() != 1)
  {
    const long long n = launch_tmm_tiled<T>(Na, Nb, Tr, I0, I1, I2, dB, dA, batch, stream);
    if (n < 0)
      return (int)-n;
    done = (size_t)n;
  }
  if (done == batch)
    return FEMTC_OK;
  const size_t nm = (size_t)Na * Nb, nb3 = (size_t)Nb * Nb * Nb, na3 = (size_t)Na * Na * Na;
  struct Tables // stream-ordered, handed back on every exit path
  {
    T* p = nullptr;
    cudaStream_t s;
    ~Tables()
    {
      if (p && cudaFreeAsync(p, s) != cudaSuccess)
        (void)cudaGetLastError();
    }
  } t;
  t.s = stream;
  FEMTC_CUDA(cudaMallocAsync(&t.p, 3 * nm * sizeof(T), stream));
  T* dI = t.p;
  FEMTC_CUDA(cudaMemcpyAsync(dI, I0, nm * sizeof(T), cudaMemcpyHostToDevice, stream));
  FEMTC_CUDA(cudaMemcpyAsync(dI + nm, I1, nm * sizeof(T), cudaMemcpyHostToDevice, stream));
  FEMTC_CUDA(cudaMemcpyAsync(dI + 2 * nm, I2, nm * sizeof(T), cudaMemcpyHostToDevice, stream));
  return launch_apply_contractions<T>(Na, Nb, Tr, dI, dI + nm, dI + 2 * nm, dB + done * nb3, dA + done * na3,
                                      batch - done, stream);
}

struct DevBuf // a call-private device buffer (batches of tensors: too large for the scratch)
{
  void* p = nullptr;
  ~DevBuf()
  {
    if (p)
      cudaFree(p);
  }
  int alloc(size_t bytes)
  {
    FEMTC_CUDA(cudaMalloc(&p, std::max<size_t>(bytes, 16)));
    return FEMTC_OK;
  }
  int upload(const void* h, size_t bytes)
  {
    int rc = alloc(bytes);
    if (rc)
      return rc;
    FEMTC_CUDA(cudaMemcpy(p, h, bytes, cudaMemcpyHostToDevice));
    return FEMTC_OK;
  }
};

template <typename T>
static int apply_contractions_host(int Na, int Nb, int Tr, const T* I0, const T* I1, const T* I2,
                                   const T* B, T* A, size_t batch)
{
  if (Na < 1 || Nb < 1 || Na > 64 || Nb > 64 || !I0 || !I1 || !I2 || (batch && (!A || !B)))
  {
    set_error("bad apply_contractions shape or null pointer");
    return FEMTC_ERR_ARG;
  }
  int rc = require_device();
  if (rc)
    return rc;
  if (batch == 0)
    return FEMTC_OK;
  const size_t nb3 = (size_t)Nb * Nb * Nb, na3 = (size_t)Na * Na * Na;
  DevBuf dB, dA;
  if ((rc = dB.upload(B, sizeof(T) * nb3 * batch)) || (rc = dA.alloc(sizeof(T) * na3 * batch)))
    return rc;
  if ((rc = apply_contractions_device<T>(Na, Nb, Tr, I0, I1, I2, (const T*)dB.p, (T*)dA.p, batch, nullptr)))
    return rc;
  FEMTC_CUDA(cudaMemcpy(A, dA.p, sizeof(T) * na3 * batch, cudaMemcpyDeviceToHost));
  return FEMTC_OK;
}

// ---- operator handle: geometry resident in HBM ------------------------------------------------------
// In a matrix-free solver detJ / G are set-up data: the reference's benches apply the operator `nreps`
// times to the same arrays (bench_mass.cpp:95-105, bench_stiffness.cpp:95-105).  A handle uploads the
// geometry once; an application then moves only U in and W out (2 of the 3 / 8 words per DoF).
struct Operator
{
  int kind = 0;        // 0 mass, 1 stiffness
  int tsize = 8;       // sizeof(T)
  int degree = 0;
  int mode = 0;
  int device = 0;
  size_t num_cells = 0;
  size_t phi_len = 0;
  std::vector<unsigned char> phi; // host copy of the tables (they travel as kernel parameters)
  void* dGeom = nullptr;
  size_t geom_bytes = 0;
};

template <typename T>
static int operator_create(Operator** out, int kind, const T* phi, size_t phi_len, const T* geom, size_t num_cells,
                           int degree, int mode)
{
  if (out)
    *out = nullptr;
  if (!out || !phi || (kind != 0 && kind != 1) || (num_cells && !geom))
  {
    set_error("femtc_operator_create: null pointer or unknown operator kind");
    return FEMTC_ERR_ARG;
  }
  if (degree < 1 || degree > 25)
  {
    set_error("degree not supported");
    return FEMTC_ERR_DEGREE;
  }
  if (kind == 1 && mode != FEMTC_STIFFNESS_REF_EXACT && mode != FEMTC_STIFFNESS_FULL)
  {
    set_error("unknown stiffness mode");
    return FEMTC_ERR_MODE;
  }
  const size_t n1 = degree + 1, n2 = n1 * n1, nd = n2 * n1;
  if (phi_len < n2)
  {
    set_error("phi shorter than (degree+1)^2");
    return FEMTC_ERR_LENGTH;
  }
  int rc = require_device();
  if (rc)
    return rc;
  Operator* op = new Operator;
  op->kind = kind, op->tsize = (int)sizeof(T), op->degree = degree, op->mode = kind ? mode : 0;
  op->num_cells = num_cells;
  op->phi_len = std::min(phi_len, 2 * n2);
  op->phi.assign(reinterpret_cast<const unsigned char*>(phi),
                 reinterpret_cast<const unsigned char*>(phi) + op->phi_len * sizeof(T));
  op->geom_bytes = num_cells * nd * (kind ? 6 : 1) * sizeof(T);
  cudaError_t e = cudaGetDevice(&op->device);
  if (e == cudaSuccess && op->geom_bytes)
    e = cudaMalloc(&op->dGeom, op->geom_bytes);
  if (e == cudaSuccess && op->geom_bytes)
    e = cudaMemcpy(op->dGeom, geom, op->geom_bytes, cudaMemcpyHostToDevice);
  if (e != cudaSuccess)
  {
    if (op->dGeom)
      cudaFree(op->dGeom);
    delete op;
    return cuda_fail(e, "femtc_operator_create (geometry upload)", __FILE__, __LINE__);
  }
  *out = op;
  return FEMTC_OK;
}

template <typename T>
static int operator_apply_device_t(const Operator* op, const T* dU, T* dW, size_t first_cell, size_t n, cudaStream_t s)
{
  const T* phi = reinterpret_cast<const T*>(op->phi.data());
  const size_t n1 = op->degree + 1, nd = n1 * n1 * n1;
  const T* g = reinterpret_cast<const T*>(op->dGeom) + first_cell * nd * (op->kind ? 6 : 1);
  if (op->kind == 0)
    return mass_device<T>(phi, op->phi_len, dU, dW, g, n, op->degree, s);
  return stiffness_device<T>(phi, op->phi_len, dU, dW, g, n, op->degree, op->mode, s);
}

template <typename T>
static int operator_apply_host_t(const Operator* op, const T* U, T* W)
{
  NvtxRange nvtx(op->kind == 0 ? region<T>("massd:resident", "massf:resident")
                               : region<T>("stiffnessd:resident", "stiffnessf:resident"));
  const size_t n1 = op->degree + 1, nd = n1 * n1 * n1;
  return host_pipeline<T>(U, W, (const T*)nullptr, op->kind ? 6 : 1, op->num_cells, nd,
                          [&](const T* dU, T* dW, const T* dG, size_t n, cudaStream_t s)
                          {
                            const T* phi = reinterpret_cast<const T*>(op->phi.data());
                            if (op->kind == 0)
                              return mass_device<T>(phi, op->phi_len, dU, dW, dG, n, op->degree, s);
                            return stiffness_device<T>(phi, op->phi_len, dU, dW, dG, n, op->degree, op->mode, s);
                          },
                          reinterpret_cast<const T*>(op->dGeom));
}

static int operator_check(const Operator* op, const void* U, const void* W)
{
  if (!op || (op->num_cells && (!U || !W)))
  {
    set_error("femtc_operator_apply: null handle or pointer");
    return FEMTC_ERR_ARG;
  }
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev != op->device)
  {
    set_error("femtc_operator_apply: the handle belongs to another device than the calling thread's current one");
    return FEMTC_ERR_ARG;
  }
  return FEMTC_OK;
}

template <typename T>
static int kernel_info_t(int op, int degree, int mode, size_t num_cells)
{
  static const T tables[2 * 26 * 26] = {};
  // never dereferenced on the host and never launched: the dispatcher only looks at alignment
  T* fake = reinterpret_cast<T*>(uintptr_t(1) << 20);
  const size_t n2 = (size_t)(degree + 1) * (degree + 1);
  if (op == 0)
    return mass_device<T>(tables, 2 * n2, fake, fake, fake, num_cells, degree, nullptr);
  return stiffness_device<T>(tables, 2 * n2, fake, fake, fake, num_cells, degree, mode, nullptr);
}

} // namespace femtc

// ================================ extern "C" =====================================================
using namespace femtc;

extern "C"
{

int femtc_mass_apply_f64(const double* phi, size_t phi_len, const double* U, double* W, const double* detJ,
                         size_t num_cells, int degree, int order, int Mb, int Nb)
{
  return mass_host<double>(phi, phi_len, U, W, detJ, num_cells, degree, order, Mb, Nb);
}
int femtc_mass_apply_f32(const float* phi, size_t phi_len, const float* U, float* W, const float* detJ,
                         size_t num_cells, int degree, int order, int Mb, int Nb)
{
  return mass_host<float>(phi, phi_len, U, W, detJ, num_cells, degree, order, Mb, Nb);
}
int femtc_stiffness_apply_f64(const double* phi, size_t phi_len, const double* U, double* W, const double* G,
                              size_t num_cells, int degree, int order, int Mb, int Nb, int mode)
{
  return stiffness_host<double>(phi, phi_len, U, W, G, num_cells, degree, order, Mb, Nb, mode);
}
int femtc_stiffness_apply_f32(const float* phi, size_t phi_len, const float* U, float* W, const float* G,
                              size_t num_cells, int degree, int order, int Mb, int Nb, int mode)
{
  return stiffness_host<float>(phi, phi_len, U, W, G, num_cells, degree, order, Mb, Nb, mode);
}

int femtc_mass_apply_device_f64(const double* phi, size_t phi_len, const double* dU, double* dW,
                                const double* ddetJ, size_t num_cells, int degree, void* stream)
{
  return mass_device<double>(phi, phi_len, dU, dW, ddetJ, num_cells, degree, (cudaStream_t)stream);
}
int femtc_mass_apply_device_f32(const float* phi, size_t phi_len, const float* dU, float* dW,
                                const float* ddetJ, size_t num_cells, int degree, void* stream)
{
  return mass_device<float>(phi, phi_len, dU, dW, ddetJ, num_cells, degree, (cudaStream_t)stream);
}
int femtc_stiffness_apply_device_f64(const double* phi, size_t phi_len, const double* dU, double* dW,
                                     const double* dG, size_t num_cells, int degree, int mode, void* stream)
{
  return stiffness_device<double>(phi, phi_len, dU, dW, dG, num_cells, degree, mode, (cudaStream_t)stream);
}
int femtc_stiffness_apply_device_f32(const float* phi, size_t phi_len, const float* dU, float* dW,
                                     const float* dG, size_t num_cells, int degree, int mode, void* stream)
{
  return stiffness_device<float>(phi, phi_len, dU, dW, dG, num_cells, degree, mode, (cudaStream_t)stream);
}

int femtc_mass_apply_device_ex_f64(const double* phi, size_t phi_len, const double* dU, double* dW, const double* ddetJ,
                                   size_t num_cells, int degree, double* d_checksum4, void* stream)
{
  return mass_device<double>(phi, phi_len, dU, dW, ddetJ, num_cells, degree, (cudaStream_t)stream, d_checksum4);
}
int femtc_mass_apply_device_ex_f32(const float* phi, size_t phi_len, const float* dU, float* dW, const float* ddetJ,
                                   size_t num_cells, int degree, double* d_checksum4, void* stream)
{
  return mass_device<float>(phi, phi_len, dU, dW, ddetJ, num_cells, degree, (cudaStream_t)stream, d_checksum4);
}
int femtc_stiffness_apply_device_ex_f64(const double* phi, size_t phi_len, const double* dU, double* dW, const double* dG,
                                        size_t num_cells, int degree, int mode, double* d_checksum4, void* stream)
{
  return stiffness_device<double>(phi, phi_len, dU, dW, dG, num_cells, degree, mode, (cudaStream_t)stream, d_checksum4);
}
int femtc_stiffness_apply_device_ex_f32(const float* phi, size_t phi_len, const float* dU, float* dW, const float* dG,
                                        size_t num_cells, int degree, int mode, double* d_checksum4, void* stream)
{
  return stiffness_device<float>(phi, phi_len, dU, dW, dG, num_cells, degree, mode, (cudaStream_t)stream, d_checksum4);
}

int femtc_gemm_blocked_f64(int k, int m, int n, int order, int MB, int NB, const double* a, const double* b, double* c)
{
  return gemm_blocked_host<double>(k, m, n, order, MB, NB, a, b, c);
}
int femtc_gemm_blocked_f32(int k, int m, int n, int order, int MB, int NB, const float* a, const float* b, float* c)
{
  return gemm_blocked_host<float>(k, m, n, order, MB, NB, a, b, c);
}
int femtc_gemm_micro_f64(int k, int m, int nc, int lda, int ldb, int ldc, int order, const double* a,
                         const double* b, double* c)
{
  return gemm_micro_host<double>(k, m, nc, lda, ldb, ldc, order, a, b, c);
}
int femtc_gemm_micro_f32(int k, int m, int nc, int lda, int ldb, int ldc, int order, const float* a,
                         const float* b, float* c)
{
  return gemm_micro_host<float>(k, m, nc, lda, ldb, ldc, order, a, b, c);
}
int femtc_gemm_strided_acc_f64(int k, int m, int nc, int sai, int sap, int sbp, int sbj, int ldc, const double* a,
                               const double* b, double* c)
{
  return gemm_strided_host<double>(k, m, nc, sai, sap, sbp, sbj, ldc, a, b, c);
}
int femtc_gemm_strided_acc_f32(int k, int m, int nc, int sai, int sap, int sbp, int sbj, int ldc, const float* a,
                               const float* b, float* c)
{
  return gemm_strided_host<float>(k, m, nc, sai, sap, sbp, sbj, ldc, a, b, c);
}
int femtc_apply_contractions_f64(int Na, int Nb, int Tr, const double* I0, const double* I1, const double* I2,
                                 const double* B, double* A, size_t batch)
{
  return apply_contractions_host<double>(Na, Nb, Tr, I0, I1, I2, B, A, batch);
}
int femtc_apply_contractions_f32(int Na, int Nb, int Tr, const float* I0, const float* I1, const float* I2,
                                 const float* B, float* A, size_t batch)
{
  return apply_contractions_host<float>(Na, Nb, Tr, I0, I1, I2, B, A, batch);
}

struct femtc_operator; /* opaque: femtc::Operator */
int femtc_operator_create_f64(femtc_operator** op, int kind, const double* phi, size_t phi_len, const double* geom,
                              size_t num_cells, int degree, int mode)
{
  return operator_create<double>(reinterpret_cast<Operator**>(op), kind, phi, phi_len, geom, num_cells, degree, mode);
}
int femtc_operator_create_f32(femtc_operator** op, int kind, const float* phi, size_t phi_len, const float* geom,
                              size_t num_cells, int degree, int mode)
{
  return operator_create<float>(reinterpret_cast<Operator**>(op), kind, phi, phi_len, geom, num_cells, degree, mode);
}
int femtc_operator_apply(femtc_operator* h, const void* U, void* W)
{
  const Operator* op = reinterpret_cast<const Operator*>(h);
  int rc = operator_check(op, U, W);
  if (rc || op->num_cells == 0)
    return rc;
  return op->tsize == 8 ? operator_apply_host_t<double>(op, (const double*)U, (double*)W)
                        : operator_apply_host_t<float>(op, (const float*)U, (float*)W);
}
int femtc_operator_apply_device(femtc_operator* h, const void* dU, void* dW, void* stream)
{
  const Operator* op = reinterpret_cast<const Operator*>(h);
  int rc = operator_check(op, dU, dW);
  if (rc || op->num_cells == 0)
    return rc;
  return op->tsize == 8
             ? operator_apply_device_t<double>(op, (const double*)dU, (double*)dW, 0, op->num_cells, (cudaStream_t)stream)
             : operator_apply_device_t<float>(op, (const float*)dU, (float*)dW, 0, op->num_cells, (cudaStream_t)stream);
}
size_t femtc_operator_device_bytes(const femtc_operator* h)
{
  return h ? reinterpret_cast<const Operator*>(h)->geom_bytes : 0;
}
void femtc_operator_destroy(femtc_operator* h)
{
  Operator* op = reinterpret_cast<Operator*>(h);
  if (!op)
    return;
  if (op->dGeom)
    cudaFree(op->dGeom);
  delete op;
}

int femtc_apply_contractions_device_f64(int Na, int Nb, int Tr, const double* I0, const double* I1, const double* I2,
                                        const double* dB, double* dA, size_t batch, void* stream)
{
  return apply_contractions_device<double>(Na, Nb, Tr, I0, I1, I2, dB, dA, batch, (cudaStream_t)stream);
}
int femtc_apply_contractions_device_f32(int Na, int Nb, int Tr, const float* I0, const float* I1, const float* I2,
                                        const float* dB, float* dA, size_t batch, void* stream)
{
  return apply_contractions_device<float>(Na, Nb, Tr, I0, I1, I2, dB, dA, batch, (cudaStream_t)stream);
}

// The general contraction kernel with its scratch in the library's own __device__ array (generic_kernels.cu): for
// callers that must not cause a device allocation.  Same argument checks as apply_contractions_device; shapes whose
// scratch does not fit the array are refused (launch_apply_contractions_static words the message).
int femtc_apply_contractions_device_static_f64(int Na, int Nb, int Tr, const double* I0, const double* I1,
                                               const double* I2, const double* dB, double* dA, size_t batch,
                                               void* stream)
{
  if (Na < 1 || Nb < 1 || Na > 64 || Nb > 64 || !I0 || !I1 || !I2 || (batch && (!dA || !dB)))
  {
    set_error("bad apply_contractions shape or null pointer");
    return FEMTC_ERR_ARG;
  }
  int rc = require_device();
  if (rc)
    return rc;
  return launch_apply_contractions_static(Na, Nb, Tr, I0, I1, I2, dB, dA, batch, (cudaStream_t)stream);
}

int femtc_string2order(const char* name)
{
  static const char* names[6] = {"ijk", "ikj", "jik", "jki", "kij", "kji"};
  if (!name)
    return -1;
  for (int i = 0; i < 6; ++i)
    if (std::strcmp(name, names[i]) == 0)
      return i;
  return -1;
}

int femtc_checksum_device_f64(const double* dX, size_t n, double* d_out4, void* stream)
{
  int rc = require_device();
  return rc ? rc : launch_checksum<double>(dX, n, d_out4, (cudaStream_t)stream);
}
int femtc_checksum_device_f32(const float* dX, size_t n, double* d_out4, void* stream)
{
  int rc = require_device();
  return rc ? rc : launch_checksum<float>(dX, n, d_out4, (cudaStream_t)stream);
}
int femtc_fill_uniform_device_f64(double* dX, size_t n, uint64_t seed, uint64_t stream_id, uint64_t offset,
                                  int centred, void* stream)
{
  int rc = require_device();
  return rc ? rc : launch_fill_uniform<double>(dX, n, seed, stream_id, offset, centred, (cudaStream_t)stream);
}
int femtc_fill_uniform_device_f32(float* dX, size_t n, uint64_t seed, uint64_t stream_id, uint64_t offset,
                                  int centred, void* stream)
{
  int rc = require_device();
  return rc ? rc : launch_fill_uniform<float>(dX, n, seed, stream_id, offset, centred, (cudaStream_t)stream);
}

int femtc_stream_probe_device(const void* d_in, void* d_out, size_t bytes_per_stream, int reads, void* stream)
{
  if ((bytes_per_stream && (!d_in || !d_out)) || ((uintptr_t)d_in & 15) || ((uintptr_t)d_out & 15) || (bytes_per_stream & 15))
  {
    set_error("stream probe: 16-byte aligned device pointers and a multiple of 16 bytes per stream required");
    return FEMTC_ERR_ARG;
  }
  int rc = require_device();
  return rc ? rc : launch_stream_probe(d_in, d_out, bytes_per_stream, reads, (cudaStream_t)stream);
}

int femtc_device_count(void)
{
  std::call_once(g_probe_once, probe_devices);
  return g_usable_devices;
}
int femtc_set_device(int device)
{
  int rc = require_device();
  if (rc)
    return rc;
  if (device < 0 || device >= g_usable_devices)
  {
    set_error("device index out of range");
    return FEMTC_ERR_ARG;
  }
  FEMTC_CUDA(cudaSetDevice(device));
  return FEMTC_OK;
}
int femtc_get_device(void)
{
  int d = -1;
  if (require_device() || cudaGetDevice(&d) != cudaSuccess)
    return -1;
  return d;
}
const char* femtc_last_error(void) { return t_last_error.c_str(); }
const char* femtc_version(void) { return "femtc_b200 0.1 (sm_100a)"; }

void* femtc_host_alloc(size_t bytes)
{
  void* p = nullptr;
  if (require_device())
    return nullptr;
  if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess)
  {
    (void)cudaGetLastError();
    set_error("cudaHostAlloc failed");
    return nullptr;
  }
  return p;
}
void femtc_host_free(void* p)
{
  if (p)
    cudaFreeHost(p);
}
void femtc_release_cache(void)
{
  unregister_all();
  drop_cached_geometry();
  int cur = 0;
  if (g_usable_devices <= 0 || cudaGetDevice(&cur) != cudaSuccess)
    return;
  for (int d = 0; d < g_usable_devices; ++d)
  {
    DeviceCache& c = g_cache[d];
    std::lock_guard<std::mutex> lock(c.mu);
    if (!c.streams_ready && !c.scratch)
      continue;
    cudaSetDevice(d);
    if (c.streams_ready)
      for (auto& s : c.slots)
    {
      cudaStreamSynchronize(s.stream);
      cudaFree(s.dU);
      cudaFree(s.dG);
      cudaFree(s.dW);
      s.dU = s.dG = s.dW = nullptr;
      s.capU = s.capG = s.capW = 0;
    }
    cudaFree(c.scratch);
    c.scratch = nullptr;
    c.scratch_cap = 0;
  }
  cudaSetDevice(cur);
}

int femtc_set_option(const char* name, long long value)
{
  if (!name)
    return FEMTC_ERR_ARG;
  if (!std::strcmp(name, "kernel") && value >= 0 && value <= 2)
    g_opt_kernel = value;
  else if (!std::strcmp(name, "host_chunk_mb") && value >= 1 && value <= 4096)
    g_opt_host_chunk_mb = value;
  else if (!std::strcmp(name, "host_register") && (value == 0 || value == 1))
    g_opt_host_register = value;
  else if (!std::strcmp(name, "host_devices") && value >= 1 && value <= 64)
    g_opt_host_devices = value;
  else if (!std::strcmp(name, "host_devices_oversubscribe") && (value == 0 || value == 1))
    g_opt_host_oversubscribe = value;
  else if (!std::strcmp(name, "geom_cache") && (value == 0 || value == 1))
    g_opt_geom_cache = value;
  else if (!std::strcmp(name, "variant"))
    g_opt_variant = value;
  else if (!std::strcmp(name, "tc") && value >= -1 && value <= 1)
    g_opt_tc = value;
  else
  {
    set_error("unknown option or value out of range");
    return FEMTC_ERR_ARG;
  }
  return FEMTC_OK;
}
long long femtc_get_option(const char* name)
{
  if (!name)
    return -1;
  if (!std::strcmp(name, "kernel"))
    return g_opt_kernel;
  if (!std::strcmp(name, "host_chunk_mb"))
    return g_opt_host_chunk_mb;
  if (!std::strcmp(name, "host_register"))
    return g_opt_host_register;
  if (!std::strcmp(name, "host_devices"))
    return g_opt_host_devices;
  if (!std::strcmp(name, "host_devices_oversubscribe"))
    return g_opt_host_oversubscribe;
  if (!std::strcmp(name, "geom_cache"))
    return g_opt_geom_cache;
  if (!std::strcmp(name, "geom_cache_uploads"))
    return g_geom_uploads;
  if (!std::strcmp(name, "variant"))
    return g_opt_variant;
  if (!std::strcmp(name, "tc"))
    return g_opt_tc;
  if (!std::strcmp(name, "launches"))
    return g_launches;
  return -1;
}

int femtc_kernel_info(int op, int dtype_bytes, int degree, int mode, long long info[8])
{
  if (!info || (op != 0 && op != 1) || (dtype_bytes != 4 && dtype_bytes != 8))
    return FEMTC_ERR_ARG;
  for (int i = 0; i < 8; ++i)
    info[i] = 0;
  if (degree < 1 || degree > 25)
    return FEMTC_ERR_DEGREE;
  int rc = require_device();
  if (rc)
    return rc;
  // walk the real dispatcher in dry-run mode for a 1 Mi-cell aligned batch: the launchers plan and report, nothing is launched
  t_record = LaunchRecord{};
  t_dry_run = true;
  rc = dtype_bytes == 8 ? kernel_info_t<double>(op, degree, mode, size_t(1) << 20)
                        : kernel_info_t<float>(op, degree, mode, size_t(1) << 20);
  t_dry_run = false;
  if (rc)
    return rc;
  const LaunchRecord& r = t_record;
  info[0] = r.family, info[1] = r.cells_per_tile, info[2] = r.threads, info[3] = r.smem;
  info[4] = r.blocks_per_sm, info[5] = r.regs, info[6] = r.grid, info[7] = r.detail;
  return FEMTC_OK;
}

} // extern "C"
