// jit.cu -- bin_fast_kernel compiled at run time for a fit's OWN systematic program.
//
// bin_fast_kernel runs a fit's systematics fastest when the program -- which systematic type acts
// on which register slot, in which order -- is a template parameter (measured on workload C/D:
// 0.85-0.89 of the HBM peak against 0.59 with run-time dispatch; profiles/r02_bin_variants.jsonl).
// The build instantiates the programs of the BASELINE workloads; every other fit gets the same
// kernel through NVRTC the first time it is planned: the translation unit is two lines (include
// bin_device.cuh, name the instantiation), compiled for sm_100a with the flags of the Makefile
// (-fmad=false: the bin arithmetic must not be contracted), loaded with the driver API and cached
// per process (a compile takes ~0.4 s, once per program, at planning time).  NVRTC and the
// driver library are dlopen'd: a box without them simply keeps the run-time dispatching kernel.
#include "common.cuh"
#include "kernels.h"

#include <cuda.h>
#include <dlfcn.h>
#include <nvrtc.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace sxmc {

namespace {

struct Api {
  bool tried = false, ok = false;
  std::string why;
  // NVRTC
  nvrtcResult (*CreateProgram)(nvrtcProgram*, const char*, const char*, int, const char* const*, const char* const*);
  nvrtcResult (*DestroyProgram)(nvrtcProgram*);
  nvrtcResult (*CompileProgram)(nvrtcProgram, int, const char* const*);
  nvrtcResult (*GetCUBINSize)(nvrtcProgram, size_t*);
  nvrtcResult (*GetCUBIN)(nvrtcProgram, char*);
  nvrtcResult (*GetProgramLogSize)(nvrtcProgram, size_t*);
  nvrtcResult (*GetProgramLog)(nvrtcProgram, char*);
  nvrtcResult (*AddNameExpression)(nvrtcProgram, const char*);
  nvrtcResult (*GetLoweredName)(nvrtcProgram, const char*, const char**);
  // driver
  CUresult (*ModuleLoadData)(CUmodule*, const void*);
  CUresult (*ModuleGetFunction)(CUfunction*, CUmodule, const char*);
  CUresult (*FuncSetAttribute)(CUfunction, CUfunction_attribute, int);
  CUresult (*LaunchKernelEx)(const CUlaunchConfig*, CUfunction, void**, void**);
  CUresult (*GetErrorString)(CUresult, const char**);
};
Api g_api;
std::mutex g_mu;

template <typename F>
bool sym(void* lib, const char* name, F& out) {
  out = (F)dlsym(lib, name);
  return out != nullptr;
}

bool load_nvrtc(Api& a) {
  const char* names[] = {"libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so.12"};
  void* lib = nullptr;
  for (const char* n : names) { lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
  if (!lib) { a.why = std::string("cannot load libnvrtc: ") + dlerror(); return false; }
  return sym(lib, "nvrtcCreateProgram", a.CreateProgram) && sym(lib, "nvrtcDestroyProgram", a.DestroyProgram) &&
         sym(lib, "nvrtcCompileProgram", a.CompileProgram) && sym(lib, "nvrtcGetCUBINSize", a.GetCUBINSize) &&
         sym(lib, "nvrtcGetCUBIN", a.GetCUBIN) && sym(lib, "nvrtcGetProgramLogSize", a.GetProgramLogSize) &&
         sym(lib, "nvrtcGetProgramLog", a.GetProgramLog) && sym(lib, "nvrtcAddNameExpression", a.AddNameExpression) &&
         sym(lib, "nvrtcGetLoweredName", a.GetLoweredName);
}

bool load_driver(Api& a) {
  void* lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) { a.why = std::string("cannot load libcuda: ") + dlerror(); return false; }
  return sym(lib, "cuModuleLoadData", a.ModuleLoadData) && sym(lib, "cuModuleGetFunction", a.ModuleGetFunction) &&
         sym(lib, "cuFuncSetAttribute", a.FuncSetAttribute) && sym(lib, "cuLaunchKernelEx", a.LaunchKernelEx) &&
         sym(lib, "cuGetErrorString", a.GetErrorString);
}

// directory of the kernel sources: $SXMC_B200_CSRC, else csrc/ next to this shared library
std::string source_dir() {
  if (const char* e = std::getenv("SXMC_B200_CSRC")) return e;
  Dl_info info;
  if (dladdr((const void*)&source_dir, &info) && info.dli_fname) {
    std::string p = info.dli_fname;
    const size_t k = p.rfind('/');
    return (k == std::string::npos ? std::string(".") : p.substr(0, k)) + "/csrc";
  }
  return "sxmc_b200/csrc";
}

struct Entry {
  std::vector<char> cubin;
  std::string lowered;
  CUfunction fn = nullptr;
  size_t smem_set = 0;
  bool failed = false;
};
std::map<std::string, Entry> g_cache;

std::string instantiation(int nobs, int ns, bool smem, int nops, unsigned codes) {
  char b[160];
  std::snprintf(b, sizeof b, "sxmc::bin_fast_kernel<%d, %d, %s, %d, %uu>", nobs, ns, smem ? "true" : "false", nops, codes);
  return b;
}

}  // namespace

// NVRTC-compile the instantiation (no GPU needed).  Returns SXMC_OK and the cubin in `e`.
static int jit_compile(Api& a, int nobs, int ns, bool smem, int nops, unsigned codes, Entry& e) {
  const std::string dir = source_dir();
  const std::string name = instantiation(nobs, ns, smem, nops, codes);
  const std::string src = "#include \"bin_device.cuh\"\ntemplate __global__ void " + name + "(const __grid_constant__ sxmc::FastParams);\n";
  nvrtcProgram prog = nullptr;
  if (a.CreateProgram(&prog, src.c_str(), "sxmc_jit.cu", 0, nullptr, nullptr) != NVRTC_SUCCESS) {
    set_error("nvrtcCreateProgram failed");
    return SXMC_ECUDA;
  }
  a.AddNameExpression(prog, name.c_str());
  const std::string inc = "-I" + dir;
  const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-fmad=false", "-lineinfo", "-default-device",
                        inc.c_str(), "-I/usr/local/cuda/include"};
  const nvrtcResult rc = a.CompileProgram(prog, (int)(sizeof opts / sizeof opts[0]), opts);
  if (rc != NVRTC_SUCCESS) {
    size_t n = 0;
    a.GetProgramLogSize(prog, &n);
    std::string log(n, ' ');
    if (n) a.GetProgramLog(prog, &log[0]);
    a.DestroyProgram(&prog);
    set_error("NVRTC could not compile " + name + " from " + dir + ": " + log.substr(0, 1500));
    return SXMC_ECUDA;
  }
  const char* lowered = nullptr;
  a.GetLoweredName(prog, name.c_str(), &lowered);
  e.lowered = lowered ? lowered : "";
  size_t n = 0;
  a.GetCUBINSize(prog, &n);
  e.cubin.resize(n);
  a.GetCUBIN(prog, e.cubin.data());
  a.DestroyProgram(&prog);
  if (e.lowered.empty() || e.cubin.empty()) { set_error("NVRTC produced no cubin for " + name); return SXMC_ECUDA; }
  return SXMC_OK;
}

static bool api_ready(bool need_driver) {
  if (!g_api.tried) {
    g_api.tried = true;
    g_api.ok = load_nvrtc(g_api);
  }
  if (!g_api.ok) return false;
  if (need_driver && !g_api.ModuleLoadData && !load_driver(g_api)) { g_api.ok = false; return false; }
  return true;
}

// Compile only (tests/test_abi.py runs this without a GPU).  *cubin_bytes_out = size of the cubin.
int jit_compile_check(int nobs, int ns, int smem, int nops, unsigned codes, size_t* cubin_bytes_out) {
  std::lock_guard<std::mutex> lock(g_mu);
  if (!api_ready(false)) { set_error("JIT unavailable: " + g_api.why); return SXMC_ECUDA; }
  Entry e;
  const int rc = jit_compile(g_api, nobs, ns, smem != 0, nops, codes, e);
  if (rc == SXMC_OK && cubin_bytes_out) *cubin_bytes_out = e.cubin.size();
  return rc;
}

// The loaded kernel for this program, or nullptr (the caller keeps the dispatching kernel).
void* jit_bin_fast_kernel(int nobs, int ns, bool smem, int nops, unsigned codes) {
  if (std::getenv("SXMC_NO_JIT")) return nullptr;
  std::lock_guard<std::mutex> lock(g_mu);
  const std::string name = instantiation(nobs, ns, smem, nops, codes);
  Entry& e = g_cache[name];
  if (e.fn) return (void*)e.fn;
  if (e.failed || !api_ready(true)) { e.failed = true; return nullptr; }
  if (e.cubin.empty() && jit_compile(g_api, nobs, ns, smem, nops, codes, e) != SXMC_OK) { e.failed = true; return nullptr; }
  cudaFree(nullptr);                                  // the primary context the runtime uses
  CUmodule mod = nullptr;
  CUresult rc = g_api.ModuleLoadData(&mod, e.cubin.data());
  if (rc == CUDA_SUCCESS) rc = g_api.ModuleGetFunction(&e.fn, mod, e.lowered.c_str());
  if (rc != CUDA_SUCCESS) {
    const char* msg = nullptr;
    g_api.GetErrorString(rc, &msg);
    set_error(std::string("loading the JIT-compiled ") + name + ": " + (msg ? msg : "?"));
    e.fn = nullptr;
    e.failed = true;
    return nullptr;
  }
  return (void*)e.fn;
}

int jit_launch_bin(void* fn, unsigned grid, size_t smem_bytes, cudaStream_t stream, bool pdl, const FastParams& P) {
  std::lock_guard<std::mutex> lock(g_mu);
  CUfunction f = (CUfunction)fn;
  CUresult rc = g_api.FuncSetAttribute(f, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, (int)smem_bytes);
  CUlaunchConfig cfg;
  std::memset(&cfg, 0, sizeof cfg);
  cfg.gridDimX = grid; cfg.gridDimY = 1; cfg.gridDimZ = 1;
  cfg.blockDimX = (unsigned)P.nwarps * 32u; cfg.blockDimY = 1; cfg.blockDimZ = 1;
  cfg.sharedMemBytes = (unsigned)smem_bytes;
  cfg.hStream = (CUstream)stream;
  CUlaunchAttribute at[1];
  std::memset(at, 0, sizeof at);
  at[0].id = CU_LAUNCH_ATTRIBUTE_PROGRAMMATIC_STREAM_SERIALIZATION;
  at[0].value.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  void* args[1] = {(void*)&P};
  if (rc == CUDA_SUCCESS) rc = g_api.LaunchKernelEx(&cfg, f, args, nullptr);
  if (rc != CUDA_SUCCESS) {
    const char* msg = nullptr;
    g_api.GetErrorString(rc, &msg);
    set_error(std::string("launching a JIT-compiled binning kernel: ") + (msg ? msg : "?"));
    return SXMC_ECUDA;
  }
  return SXMC_OK;
}

}  // namespace sxmc
