// jit.cu -- compiled-in mask constants for ANY layout, at run time.
//
// In the reference every packed_int<Integer, Bits...> gets its H / L / saturation-mask constants
// at compile time (pint.hpp:324-337, 443-456).  libpint_cuda.so carries such kernels for the layouts
// of registry_layouts.def; any other layout runs the run-time-mask kernels (same results, masks in
// the constant bank, ALU-bound on mixed widths).  With option "jit" > 0 the library instead
// specialises the SAME kernel templates (include/pint/cuda_kernels*.cuh, embedded in the library as
// text) for the layout with NVRTC -- static_layout<Integer, Bits...> exactly as the registry does --
// loads the cubin through the driver API and launches the vector body with it.  Heads, tails and
// operands that are not congruent modulo 16 stay on the run-time-mask kernels (identical results).
//
//   - libnvrtc is dlopen'ed on first use (no link-time dependency); the driver entry points come
//     from cudaGetDriverEntryPoint (no libcuda link either).
//   - One module per (device, layout, kernel family), compiled on first use, kept for the life of
//     the process; cubins are cached on disk ($PINT_CUDA_JIT_CACHE, else ~/.cache/pint_cuda) keyed
//     by a hash of the embedded headers, the generated source and the options.
//   - "jit" = 1: a layout that cannot be specialised (no NVRTC, compile error) falls back to the
//     run-time-mask kernels and pint_cuda_jit_last_error() says why; "jit" = 2: the call fails with
//     PINT_E_JIT.  Default 0: the device entry points never compile, allocate or block.
#include <dlfcn.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include <cuda.h>

#include "api_common.hpp"
#include "pint/cuda_kernels_dense.cuh"  // choose_tile, dense_tuning, kMaxMwWords (host side of the dense launcher)

namespace pint_cuda_impl {
namespace {

using pint::cuda::kernels::kNumOpPairs;
using pint::cuda::kernels::kOpPairs;

struct embedded_header {
    const char* name;
    const char* text;
};
const embedded_header kHeaders[] = {
#include "jit_headers.inc"
};
constexpr int kNumHeaders = int(sizeof(kHeaders) / sizeof(kHeaders[0]));

// ---- NVRTC, loaded on demand ----
using nvrtcProgram = struct _nvrtcProgram*;
struct nvrtc_api {
    void* handle = nullptr;
    int (*CreateProgram)(nvrtcProgram*, const char*, const char*, int, const char* const*, const char* const*) = nullptr;
    int (*DestroyProgram)(nvrtcProgram*) = nullptr;
    int (*CompileProgram)(nvrtcProgram, int, const char* const*) = nullptr;
    int (*GetProgramLogSize)(nvrtcProgram, size_t*) = nullptr;
    int (*GetProgramLog)(nvrtcProgram, char*) = nullptr;
    int (*GetCUBINSize)(nvrtcProgram, size_t*) = nullptr;
    int (*GetCUBIN)(nvrtcProgram, char*) = nullptr;
    int (*AddNameExpression)(nvrtcProgram, const char*) = nullptr;
    int (*GetLoweredName)(nvrtcProgram, const char*, const char**) = nullptr;
    int (*Version)(int*, int*) = nullptr;
    std::string include_dir;  // where <cuda/std/...> lives
};

struct driver_api {
    CUresult (*ModuleLoadData)(CUmodule*, const void*) = nullptr;
    CUresult (*ModuleUnload)(CUmodule) = nullptr;  // optional: only used to drop a module whose disk-cache entry was damaged
    CUresult (*ModuleGetFunction)(CUfunction*, CUmodule, const char*) = nullptr;
    CUresult (*ModuleGetGlobal)(CUdeviceptr*, size_t*, CUmodule, const char*) = nullptr;
    CUresult (*MemcpyDtoH)(void*, CUdeviceptr, size_t) = nullptr;
    CUresult (*LaunchKernel)(CUfunction, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, CUstream, void**, void**) = nullptr;
    CUresult (*FuncSetAttribute)(CUfunction, CUfunction_attribute, int) = nullptr;
    CUresult (*LaunchKernelEx)(const CUlaunchConfig*, CUfunction, void**, void**) = nullptr;  // optional: launches with the PDL attribute
};

std::mutex g_jit_mutex;
std::string g_last_error;
nvrtc_api g_nvrtc;
driver_api g_drv;
bool g_nvrtc_tried = false, g_drv_ok = false;

bool file_exists(const std::string& p)
{
    struct stat st;
    return ::stat(p.c_str(), &st) == 0;
}

bool load_nvrtc()
{
    if (g_nvrtc.handle) return true;
    if (g_nvrtc_tried) return false;
    g_nvrtc_tried = true;
    // Candidates in order: the override, the toolkit's own copy by absolute path, then whatever the
    // loader finds.  The 256-bit global accesses of the kernels (ld/st.global.v8) need the PTX ISA
    // of CUDA 12.9, and a process that has imported a Python CUDA stack usually already holds an
    // older libnvrtc.so.12 under the same soname: an absolute path gets the toolkit's regardless.
    std::vector<std::string> names;
    if (const char* e = std::getenv("PINT_CUDA_NVRTC")) names.push_back(e);
    for (const char* v : {"CUDA_HOME", "CUDA_PATH"})
        if (const char* e = std::getenv(v)) names.push_back(std::string(e) + "/lib64/libnvrtc.so.12");
    for (const char* n : {"/usr/local/cuda/lib64/libnvrtc.so.12", "/usr/local/cuda/targets/x86_64-linux/lib/libnvrtc.so.12", "libnvrtc.so.12", "libnvrtc.so"})
        names.push_back(n);
    void* h = nullptr;
    std::string seen;
    for (const auto& n : names) {
        void* c = dlopen(n.c_str(), RTLD_NOW | RTLD_LOCAL);
        if (!c) continue;
        int major = 0, minor = 0;
        auto ver = reinterpret_cast<int (*)(int*, int*)>(dlsym(c, "nvrtcVersion"));
        if (ver && ver(&major, &minor) == 0 && (major > 12 || (major == 12 && minor >= 9))) { h = c; break; }
        seen += " " + n + " (" + std::to_string(major) + "." + std::to_string(minor) + ")";
        dlclose(c);
    }
    if (!h) {
        g_last_error = "no libnvrtc >= 12.9 found (set PINT_CUDA_NVRTC); tried and rejected:" + (seen.empty() ? std::string(" nothing loadable") : seen);
        return false;
    }
    auto sym = [&](const char* s) { return dlsym(h, s); };
#define PINT_NVRTC_SYM(field, name) g_nvrtc.field = reinterpret_cast<decltype(g_nvrtc.field)>(sym(name))
    PINT_NVRTC_SYM(CreateProgram, "nvrtcCreateProgram");
    PINT_NVRTC_SYM(DestroyProgram, "nvrtcDestroyProgram");
    PINT_NVRTC_SYM(CompileProgram, "nvrtcCompileProgram");
    PINT_NVRTC_SYM(GetProgramLogSize, "nvrtcGetProgramLogSize");
    PINT_NVRTC_SYM(GetProgramLog, "nvrtcGetProgramLog");
    PINT_NVRTC_SYM(GetCUBINSize, "nvrtcGetCUBINSize");
    PINT_NVRTC_SYM(GetCUBIN, "nvrtcGetCUBIN");
    PINT_NVRTC_SYM(AddNameExpression, "nvrtcAddNameExpression");
    PINT_NVRTC_SYM(GetLoweredName, "nvrtcGetLoweredName");
    PINT_NVRTC_SYM(Version, "nvrtcVersion");
#undef PINT_NVRTC_SYM
    if (!g_nvrtc.CreateProgram || !g_nvrtc.CompileProgram || !g_nvrtc.GetCUBIN || !g_nvrtc.AddNameExpression || !g_nvrtc.GetLoweredName) {
        g_last_error = "libnvrtc lacks a required entry point";
        dlclose(h);
        return false;
    }
    // libcu++ (<cuda/std/...>) sits in the toolkit's include directory
    std::vector<std::string> dirs;
    if (const char* e = std::getenv("PINT_CUDA_INCLUDE")) dirs.push_back(e);
    for (const char* v : {"CUDA_HOME", "CUDA_PATH"})
        if (const char* e = std::getenv(v)) dirs.push_back(std::string(e) + "/include");
    Dl_info info;
    if (dladdr(reinterpret_cast<void*>(g_nvrtc.CreateProgram), &info) && info.dli_fname) {
        std::string p = info.dli_fname;  // .../lib64/libnvrtc.so.12 or .../targets/x86_64-linux/lib/libnvrtc.so.12
        const size_t slash = p.rfind('/');
        if (slash != std::string::npos) {
            const std::string libdir = p.substr(0, slash);
            dirs.push_back(libdir + "/../include");
        }
    }
    dirs.push_back("/usr/local/cuda/include");
    for (const auto& d : dirs)
        if (file_exists(d + "/cuda/std/type_traits")) { g_nvrtc.include_dir = d; break; }
    if (g_nvrtc.include_dir.empty()) {
        g_last_error = "libcu++ headers (cuda/std/type_traits) not found (set PINT_CUDA_INCLUDE)";
        dlclose(h);
        return false;
    }
    g_nvrtc.handle = h;
    return true;
}

bool load_driver()
{
    if (g_drv_ok) return true;
    auto get = [](const char* name, void** fn) {
        cudaDriverEntryPointQueryResult q;
        return cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess && *fn;
    };
    bool ok = get("cuModuleLoadData", reinterpret_cast<void**>(&g_drv.ModuleLoadData)) &&
              get("cuModuleGetFunction", reinterpret_cast<void**>(&g_drv.ModuleGetFunction)) &&
              get("cuModuleGetGlobal", reinterpret_cast<void**>(&g_drv.ModuleGetGlobal)) &&
              get("cuMemcpyDtoH", reinterpret_cast<void**>(&g_drv.MemcpyDtoH)) &&
              get("cuLaunchKernel", reinterpret_cast<void**>(&g_drv.LaunchKernel)) &&
              get("cuFuncSetAttribute", reinterpret_cast<void**>(&g_drv.FuncSetAttribute));
    if (!ok) { cudaGetLastError(); g_last_error = "driver entry points unavailable"; }
    if (ok && !get("cuModuleUnload", reinterpret_cast<void**>(&g_drv.ModuleUnload))) { cudaGetLastError(); g_drv.ModuleUnload = nullptr; }
    if (ok && !get("cuLaunchKernelEx", reinterpret_cast<void**>(&g_drv.LaunchKernelEx))) { cudaGetLastError(); g_drv.LaunchKernelEx = nullptr; }
    g_drv_ok = ok;
    return ok;
}

// One launch of a specialised kernel.  With option "pdl" the launch carries the programmatic-stream-serialization
// attribute, like launch_stream_kernel does for the pre-built kernels (every kernel template starts with pdl_prologue).
CUresult drv_launch(CUfunction f, unsigned grid, unsigned block, unsigned smem, cudaStream_t s, void** args, bool pdl)
{
    if (pdl && g_drv.LaunchKernelEx) {
        CUlaunchConfig cfg{};
        cfg.gridDimX = grid; cfg.gridDimY = 1; cfg.gridDimZ = 1;
        cfg.blockDimX = block; cfg.blockDimY = 1; cfg.blockDimZ = 1;
        cfg.sharedMemBytes = smem;
        cfg.hStream = (CUstream)s;
        CUlaunchAttribute attr[1];
        attr[0].id = CU_LAUNCH_ATTRIBUTE_PROGRAMMATIC_STREAM_SERIALIZATION;
        attr[0].value.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        return g_drv.LaunchKernelEx(&cfg, f, args, nullptr);
    }
    return g_drv.LaunchKernel(f, grid, 1, 1, block, 1, 1, smem, (CUstream)s, args, nullptr);
}

// ---- what gets compiled ----
// FAM_DENSE / FAM_DENSE_PAIR: the whole-chunk kernels of the dense bit-stream format (cuda_kernels_dense_mw.cuh) with the
// chunk masks of chunk_layout<Bits...> compiled in -- what dense_static.cu instantiates for the layouts of dense_layouts.def.
// They exist for layouts whose chunk is at most kMaxMwWords words and whose sum(Bits) is not 8 / 16 / 32 / 64.
// FAM_SCALAR / FAM_PAIRSMAP: the caller shapes of api_callers.cu -- one operand constant (26 kernels: op x operand side)
// and interleaved std::pair operands (13) -- with the functor registry_part.cu picks for a registered layout (caller_fn).
enum family : int { FAM_BINARY = 0, FAM_FUSED = 1, FAM_PAIR = 2, FAM_DENSE = 3, FAM_DENSE_PAIR = 4, FAM_SCALAR = 5, FAM_PAIRSMAP = 6, FAM_COUNT = 7 };
const char* const kFamilyName[FAM_COUNT] = {"binary", "fused", "pair", "dense", "densepair", "scalar", "pairsmap"};

uint32_t sum_bits(const layout_desc& d)
{
    uint32_t t = 0;
    for (uint32_t i = 0; i < d.nfields; ++i) t += d.bits[i];
    return t;
}

int dense_chunk_words(const layout_desc& d)
{
    const uint32_t S = sum_bits(d);
    return int(S / pint::cuda::kernels::dense::gcd_u32(S, 32));
}

bool is_dense_family(int fam) { return fam == FAM_DENSE || fam == FAM_DENSE_PAIR; }

bool family_applies(const layout_desc& d, int fam)
{
    if (!is_dense_family(fam)) return true;
    const uint32_t S = sum_bits(d);
    return S != 8 && S != 16 && S != 32 && S != 64 && dense_chunk_words(d) <= (int)pint::cuda::kernels::dense::kMaxMwWords;
}

// kernels per entry of a family: 16- and 32-byte vector forms for the padded families, one for the dense ones
int family_variants(int fam) { return fam <= FAM_PAIR ? 2 : 1; }

struct jit_kernel {
    CUfunction f128 = nullptr, f256 = nullptr;
    int unroll = 2, block = 256, vec256 = 0, heavy = 0;
};

struct jit_module {
    int state = 0;  // 0 = not tried, 1 = ready, -1 = failed
    CUmodule mod = nullptr;
    std::vector<jit_kernel> k;
};

struct jit_entry {
    layout_desc d;
    int device;
    jit_module fam[FAM_COUNT];
};

std::vector<jit_entry*> g_entries;

const char* integer_name(uint32_t eb)
{
    return eb == 1 ? "uint8_t" : eb == 2 ? "uint16_t" : eb == 4 ? "uint32_t" : "uint64_t";
}

int family_count(int fam)
{
    if (fam == FAM_PAIR || fam == FAM_DENSE_PAIR) return kNumOpPairs;
    if (fam == FAM_SCALAR) return 2 * (int)pint::cuda::OP_NUM_BINARY;  // index = scalar_first * 13 + op
    return (int)pint::cuda::OP_NUM_BINARY;
}

std::string kernel_fn(int fam, int i)
{
    char buf[96];
    if (fam == FAM_BINARY) std::snprintf(buf, sizeof buf, "pint_jit::BFn<%d>", i);
    else if (fam == FAM_FUSED) std::snprintf(buf, sizeof buf, "pint_jit::FFn<%d>", i);
    else if (fam == FAM_DENSE) std::snprintf(buf, sizeof buf, "pint_jit::DFn<%d>", i);
    else if (fam == FAM_DENSE_PAIR) std::snprintf(buf, sizeof buf, "pint_jit::DPFn<%d>", i);
    else if (fam == FAM_SCALAR) std::snprintf(buf, sizeof buf, "pint_jit::SFn<%d, %s>", i % (int)pint::cuda::OP_NUM_BINARY, i >= (int)pint::cuda::OP_NUM_BINARY ? "true" : "false");
    else if (fam == FAM_PAIRSMAP) std::snprintf(buf, sizeof buf, "pint_jit::CFn<%d>", i);
    else std::snprintf(buf, sizeof buf, "pint_jit::PFn<%d>", i);
    return buf;
}

std::string kernel_expr(int fam, int i, bool wide)
{
    const std::string fn = kernel_fn(fam, i);
    if (is_dense_family(fam))  // Mode 0: one result; Mode 1: the second result of an op pair leaves over the b tile
        return "pint::cuda::kernels::dense::dense_binary_kernel<" + fn + ", pint_jit::DGeo, " + (fam == FAM_DENSE_PAIR ? "1" : "0") + ">";
    if (fam == FAM_SCALAR) return "pint::cuda::kernels::map1_stream_kernel<" + fn + ", 4>";                              // launch_map1's U
    if (fam == FAM_PAIRSMAP) return "pint::cuda::kernels::pairs_map_kernel<pint_jit::SL::integer, " + fn + ", 2>";   // launch_pairs_map's U
    const char* base = fam == FAM_PAIR ? (wide ? "binary_pair_stream256_kernel" : "binary_pair_stream_kernel")
                                       : (wide ? "binary_stream256_kernel" : "binary_stream_kernel");
    return std::string("pint::cuda::kernels::") + base + "<" + fn + ", pint_jit::Cfg<" + fn + ">>";
}

std::string make_dense_source(const layout_desc& d, int fam)
{
    std::string s = "#include \"pint/cuda_kernels_dense.cuh\"\nnamespace pint_jit {\nusing namespace pint::cuda;\nusing namespace pint::cuda::kernels;\n";
    s += "using CL = dense::chunk_layout<";
    for (uint32_t i = 0; i < d.nfields; ++i) s += (i ? ", " : "") + std::to_string((unsigned)d.bits[i]);
    s += ">;\nusing M = static_masks<CL>;\nusing DGeo = dense::mw_geom<CL::kN>;\n"
         "template <int Op> using DFn = swar_fn<Op, M>;\n"
         "template <int P> using DPFn = pair_fn<kOpPairs[P][0], kOpPairs[P][1], M>;\n"
         "}\n";
    // sanity data for the host: chunk words, element bits (must equal what the host computes from the layout)
    s += "extern \"C\" __device__ const int pint_jit_cfg[" + std::to_string(family_count(fam)) + "][4] = {\n";
    for (int i = 0; i < family_count(fam); ++i) s += "  {pint_jit::CL::kN, int(pint_jit::CL::kS), int(sizeof(pint_jit::DGeo)), 0},\n";
    s += "};\n";
    return s;
}

std::string make_source(const layout_desc& d, int fam)
{
    if (is_dense_family(fam)) return make_dense_source(d, fam);
    std::string s = "#include \"pint/cuda_kernels.cuh\"\nnamespace pint_jit {\nusing namespace pint::cuda;\nusing namespace pint::cuda::kernels;\n";
    s += "using SL = static_layout<" + std::string(integer_name(d.elem_bytes));
    for (uint32_t i = 0; i < d.nfields; ++i) s += ", " + std::to_string((unsigned)d.bits[i]);
    s += ">;\nusing M = static_masks<SL>;\nusing F = static_fields<SL>;\n"
         "template <int Op> using BFn = swar_fn<Op, M>;\n"
         "template <int Op> using FFn = fused_fn<Op, F>;\n"
         "template <int P> using PFn = pair_fn<kOpPairs[P][0], kOpPairs[P][1], M>;\n"
         "template <int Op> using CFn = typename std::conditional<sizeof(typename M::word) == 8 && swar_fn<Op, M>::heavy, fused_fn<Op, F>, swar_fn<Op, M>>::type;\n"
         "template <int Op, bool First> using SFn = scalar_fn<CFn<Op>, First>;\n"
         "template <class Fn> using Cfg = stream_cfg_for<Fn>;\n"
         "template <class Fn> struct wide { static constexpr int value = int(Fn::vec256); };\n"
         "template <int A, int B, class P> struct wide<pair_fn<A, B, P>> { static constexpr int value = (!pair_fn<A, B, P>::heavy && sizeof(typename P::word) == 4) ? 1 : 0; };\n"
         "}\n";
    // per kernel: unroll, block, whether the library's launcher would pick 32-byte vectors, heavy
    s += "extern \"C\" __device__ const int pint_jit_cfg[" + std::to_string(family_count(fam)) + "][4] = {\n";
    for (int i = 0; i < family_count(fam); ++i) {
        const std::string fn = kernel_fn(fam, i);
        if (fam == FAM_SCALAR)  // size of the functor parameter and where its scalar word sits: the host builds the same bytes
            // k is scalar_fn's last member and as wide as the struct's alignment: it sits in the last sizeof(word) bytes
            s += "  {int(sizeof(" + fn + ")), int(sizeof(" + fn + ") - sizeof(typename " + fn + "::word)), int(sizeof(typename " + fn + "::word)), 0},\n";
        else if (fam == FAM_PAIRSMAP)
            s += "  {int(sizeof(" + fn + ")), 0, 0, 0},\n";
        else
            s += "  {pint_jit::Cfg<" + fn + ">::UNROLL, pint_jit::Cfg<" + fn + ">::BLOCK, pint_jit::wide<" + fn + ">::value, int(" + fn + "::heavy)},\n";
    }
    s += "};\n";
    return s;
}

uint64_t fnv1a(uint64_t h, const void* p, size_t n)
{
    const unsigned char* c = static_cast<const unsigned char*>(p);
    for (size_t i = 0; i < n; ++i) { h ^= c[i]; h *= 0x100000001B3ULL; }
    return h;
}

std::string cache_dir()
{
    std::string dir;
    if (const char* e = std::getenv("PINT_CUDA_JIT_CACHE")) dir = e;
    else if (const char* x = std::getenv("XDG_CACHE_HOME")) dir = std::string(x) + "/pint_cuda";
    else if (const char* h = std::getenv("HOME")) dir = std::string(h) + "/.cache/pint_cuda";
    if (dir.empty()) return dir;
    // mkdir -p, two levels are enough for the defaults
    const size_t slash = dir.rfind('/');
    if (slash != std::string::npos && slash > 0) ::mkdir(dir.substr(0, slash).c_str(), 0755);
    ::mkdir(dir.c_str(), 0755);
    return file_exists(dir) ? dir : std::string();
}

bool read_file(const std::string& path, std::vector<char>* out)
{
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out->resize(n > 0 ? size_t(n) : 0);
    const bool ok = n > 0 && std::fread(out->data(), 1, size_t(n), f) == size_t(n);
    std::fclose(f);
    return ok;
}

void write_file_atomic(const std::string& path, const std::vector<char>& data)
{
    const std::string tmp = path + ".tmp" + std::to_string((long)::getpid());
    FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) return;
    const bool ok = std::fwrite(data.data(), 1, data.size(), f) == data.size();
    std::fclose(f);
    if (ok) std::rename(tmp.c_str(), path.c_str());
    else std::remove(tmp.c_str());
}

// Compiles (or fetches from the disk cache) one family for one layout.  Lowered kernel names are
// stored beside the cubin so that a cache hit needs no NVRTC at all.
bool build_cubin(const layout_desc& d, int fam, int cc_major, int cc_minor, std::vector<char>* cubin, std::vector<std::string>* lowered,
                 bool ignore_cache = false, bool* from_cache = nullptr)
{
    if (from_cache) *from_cache = false;
    const std::string src = make_source(d, fam);
    char arch[48];
    std::snprintf(arch, sizeof arch, "-arch=sm_%d%d%s", cc_major, cc_minor, cc_major >= 9 ? "a" : "");
    uint64_t h = 0xCBF29CE484222325ULL;
    for (int i = 0; i < kNumHeaders; ++i) h = fnv1a(h, kHeaders[i].text, std::strlen(kHeaders[i].text));
    h = fnv1a(h, src.data(), src.size());
    h = fnv1a(h, arch, std::strlen(arch));
    const std::string dir = cache_dir();
    char key[40];
    std::snprintf(key, sizeof key, "%016llx", (unsigned long long)h);
    const std::string base = dir.empty() ? std::string() : dir + "/" + key + "_" + kFamilyName[fam];
    const int nk = family_count(fam);
    if (!base.empty() && !ignore_cache) {
        std::vector<char> names;
        if (read_file(base + ".cubin", cubin) && read_file(base + ".names", &names)) {
            // first line: "#cubin <bytes> <fnv1a of the image>", so that a truncated or foreign file never reaches the driver
            lowered->clear();
            std::string cur, head;
            bool first = true;
            for (char c : names) {
                if (c == '\n') {
                    if (first) { head = cur; first = false; }
                    else lowered->push_back(cur);
                    cur.clear();
                } else cur.push_back(c);
            }
            char want[80];
            std::snprintf(want, sizeof want, "#cubin %zu %016llx", cubin->size(),
                          (unsigned long long)fnv1a(0xCBF29CE484222325ULL, cubin->data(), cubin->size()));
            if (head == want && (int)lowered->size() == family_variants(fam) * nk) {
                if (from_cache) *from_cache = true;
                return true;
            }
        }
    }
    if (!load_nvrtc()) return false;
    std::vector<const char*> hdr_text, hdr_name;
    std::vector<std::string> alt;  // every header under "pint/<name>" and "<name>" (the headers include each other by bare name)
    alt.reserve(kNumHeaders);
    for (int i = 0; i < kNumHeaders; ++i) {
        hdr_text.push_back(kHeaders[i].text);
        hdr_name.push_back(kHeaders[i].name);
        alt.push_back(std::string("pint/") + kHeaders[i].name);
    }
    for (int i = 0; i < kNumHeaders; ++i) {
        hdr_text.push_back(kHeaders[i].text);
        hdr_name.push_back(alt[i].c_str());
    }
    nvrtcProgram prog = nullptr;
    if (g_nvrtc.CreateProgram(&prog, src.c_str(), "pint_jit.cu", (int)hdr_text.size(), hdr_text.data(), hdr_name.data()) != 0) {
        g_last_error = "nvrtcCreateProgram failed";
        return false;
    }
    std::vector<std::string> exprs;
    for (int i = 0; i < nk; ++i) {
        exprs.push_back(kernel_expr(fam, i, false));
        if (family_variants(fam) == 2) exprs.push_back(kernel_expr(fam, i, true));
    }
    for (const auto& e : exprs) g_nvrtc.AddNameExpression(prog, e.c_str());
    const std::string inc = "-I" + g_nvrtc.include_dir;
    const char* opts[] = {"-std=c++17", arch, "-default-device", "-lineinfo", inc.c_str()};
    const int rc = g_nvrtc.CompileProgram(prog, 5, opts);
    if (rc != 0) {
        size_t n = 0;
        std::string log;
        if (g_nvrtc.GetProgramLogSize && g_nvrtc.GetProgramLogSize(prog, &n) == 0 && n > 1) {
            log.resize(n);
            g_nvrtc.GetProgramLog(prog, &log[0]);
        }
        g_last_error = "nvrtcCompileProgram failed (" + std::to_string(rc) + "): " + log.substr(0, 2000);
        g_nvrtc.DestroyProgram(&prog);
        return false;
    }
    lowered->clear();
    for (const auto& e : exprs) {
        const char* low = nullptr;
        if (g_nvrtc.GetLoweredName(prog, e.c_str(), &low) != 0 || !low) {
            g_last_error = "nvrtcGetLoweredName failed for " + e;
            g_nvrtc.DestroyProgram(&prog);
            return false;
        }
        lowered->push_back(low);
    }
    size_t sz = 0;
    if (g_nvrtc.GetCUBINSize(prog, &sz) != 0 || sz == 0) {
        g_last_error = "nvrtcGetCUBINSize failed";
        g_nvrtc.DestroyProgram(&prog);
        return false;
    }
    cubin->resize(sz);
    g_nvrtc.GetCUBIN(prog, cubin->data());
    g_nvrtc.DestroyProgram(&prog);
    if (!base.empty()) {
        std::vector<char> names;
        char head[80];
        const int hn = std::snprintf(head, sizeof head, "#cubin %zu %016llx\n", cubin->size(),
                                     (unsigned long long)fnv1a(0xCBF29CE484222325ULL, cubin->data(), cubin->size()));
        names.insert(names.end(), head, head + hn);
        for (const auto& l : *lowered) { names.insert(names.end(), l.begin(), l.end()); names.push_back('\n'); }
        write_file_atomic(base + ".names", names);
        write_file_atomic(base + ".cubin", *cubin);
    }
    return true;
}

// caller holds g_jit_mutex
jit_module* get_module(const layout_desc& d, int fam)
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    jit_entry* e = nullptr;
    for (jit_entry* x : g_entries)
        if (x->device == dev && same_layout(x->d, d)) { e = x; break; }
    if (!e) {
        e = new jit_entry{};
        e->d = d;
        e->device = dev;
        g_entries.push_back(e);
    }
    jit_module& m = e->fam[fam];
    if (m.state != 0) return m.state > 0 ? &m : nullptr;
    m.state = -1;
    if (!load_driver()) return nullptr;
    cudaFree(nullptr);  // the primary context must exist and be current before the module is loaded
    int major = 0, minor = 0;
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
    std::vector<char> cubin;
    std::vector<std::string> lowered;
    const int nk = family_count(fam);
    auto load = [&]() -> bool {
        // a file that is not a loadable image: the driver parses it, so refuse the obviously truncated ones first
        if (cubin.size() < 64 || std::memcmp(cubin.data(), "\177ELF", 4) != 0) { g_last_error = "not a cubin"; return false; }
        if (g_drv.ModuleLoadData(&m.mod, cubin.data()) != CUDA_SUCCESS) { g_last_error = "cuModuleLoadData failed"; m.mod = nullptr; return false; }
        m.k.assign(nk, jit_kernel{});
        std::vector<int> cfg(size_t(nk) * 4);
        CUdeviceptr p = 0;
        size_t bytes = 0;
        if (g_drv.ModuleGetGlobal(&p, &bytes, m.mod, "pint_jit_cfg") != CUDA_SUCCESS || bytes != cfg.size() * sizeof(int) ||
            g_drv.MemcpyDtoH(cfg.data(), p, bytes) != CUDA_SUCCESS) {
            g_last_error = "pint_jit_cfg missing from the module";
            return false;
        }
        for (int i = 0; i < nk; ++i) {
            jit_kernel& k = m.k[i];
            const int nv = family_variants(fam);
            if (g_drv.ModuleGetFunction(&k.f128, m.mod, lowered[nv * i].c_str()) != CUDA_SUCCESS ||
                (nv == 2 && g_drv.ModuleGetFunction(&k.f256, m.mod, lowered[nv * i + 1].c_str()) != CUDA_SUCCESS)) {
                g_last_error = "cuModuleGetFunction failed for " + lowered[nv * i];
                return false;
            }
            k.unroll = cfg[4 * i]; k.block = cfg[4 * i + 1]; k.vec256 = cfg[4 * i + 2]; k.heavy = cfg[4 * i + 3];
        }
        return true;
    };
    bool from_cache = false;
    if (!build_cubin(d, fam, major, minor, &cubin, &lowered, false, &from_cache)) return nullptr;
    if (!load()) {
        // a damaged disk-cache entry (truncated file, another toolkit's image): compile afresh once, which rewrites it
        if (!from_cache) return nullptr;
        if (m.mod && g_drv.ModuleUnload) g_drv.ModuleUnload(m.mod);
        m.mod = nullptr;
        if (!build_cubin(d, fam, major, minor, &cubin, &lowered, true) || !load()) return nullptr;
    }
    m.state = 1;
    return &m;
}

jit_module* lookup(const layout_desc& d, int fam)
{
    std::lock_guard<std::mutex> lock(g_jit_mutex);
    return get_module(d, fam);
}

using pint::cuda::kernels::count_launch;
using pint::cuda::kernels::grid_for;
using pint::cuda::kernels::split_aligned;

// the vector body of [0, n) with the specialised kernel; `rest(first, count)` runs heads and tails
template <class Rest>
int launch_body(const jit_kernel& k, uint32_t elem_bytes, const void* const* ptrs, int nptrs, size_t n, cudaStream_t s,
                const launch_geom& g, Rest rest, bool* handled)
{
    size_t head = 0, nvec = 0, tail = 0;
    bool wide = (g.vec256 == 2 || (g.vec256 == 1 && k.vec256)) && split_aligned(elem_bytes, n, ptrs, nptrs, &head, &nvec, &tail, 32) && nvec;
    if (!wide && !(split_aligned(elem_bytes, n, ptrs, nptrs, &head, &nvec, &tail, 16) && nvec)) {
        *handled = false;  // not congruent modulo 16 (or shorter than a vector): the run-time-mask kernels take all of it
        return PINT_OK;
    }
    *handled = true;
    if (int rc = rest(size_t(0), head)) return rc;
    const size_t per_thread = wide ? size_t(k.unroll / 2 > 0 ? k.unroll / 2 : 1) : size_t(k.unroll);
    const size_t tile = per_thread * size_t(k.block);
    const size_t ntiles = (nvec + tile - 1) / tile;
    const void* vp[4];
    void* args[6];
    char empty_fn = 0;  // the functor of a compiled-in layout is an empty class: one byte of parameter space
    for (int i = 0; i < nptrs; ++i) {
        vp[i] = static_cast<const char*>(ptrs[i]) + head * elem_bytes;
        args[i] = &vp[i];
    }
    args[nptrs] = &nvec;
    args[nptrs + 1] = &empty_fn;
    const CUresult cr = drv_launch(wide ? k.f256 : k.f128, grid_for(ntiles, g), (unsigned)k.block, 0, s, args, g.pdl != 0);
    if (cr != CUDA_SUCCESS) return (int)cudaErrorLaunchFailure;
    count_launch(g);
    return rest(n - tail, tail);
}

}  // namespace

int jit_binary(int op, bool fused, const pint_layout* layout, const layout_desc& d, const void* a, const void* b, void* out,
               size_t n, cudaStream_t s, const launch_geom& g, bool* handled)
{
    *handled = false;
    jit_module* m = lookup(d, fused ? FAM_FUSED : FAM_BINARY);
    if (!m) return PINT_E_JIT;
    const void* ptrs[3] = {a, b, out};
    auto rest = [&](size_t first, size_t count) -> int {
        if (!count) return PINT_OK;
        const size_t off = first * d.elem_bytes;
        return (int)generic_binary(op, fused, d, (const char*)a + off, (const char*)b + off, (char*)out + off, count, s, g);
    };
    (void)layout;
    return launch_body(m->k[op], d.elem_bytes, ptrs, 3, n, s, g, rest, handled);
}

int jit_binary_pair(int pair, const layout_desc& d, const void* a, const void* b, void* out0, void* out1, size_t n,
                    cudaStream_t s, const launch_geom& g, bool* handled)
{
    *handled = false;
    jit_module* m = lookup(d, FAM_PAIR);
    if (!m) return PINT_E_JIT;
    const void* ptrs[4] = {a, b, out0, out1};
    auto rest = [&](size_t first, size_t count) -> int {
        if (!count) return PINT_OK;
        const size_t off = first * d.elem_bytes;
        return (int)generic_binary_pair(pair, d, (const char*)a + off, (const char*)b + off, (char*)out0 + off, (char*)out1 + off, count, s, g);
    };
    return launch_body(m->k[pair], d.elem_bytes, ptrs, 4, n, s, g, rest, handled);
}

// Dense bit streams: op >= 0 runs one op (out1 unused), pair >= 0 one of kOpPairs with both results.  The launch
// geometry is launch_dense_binary's (cuda_kernels_dense.cuh): same tile chooser, same shared-memory size, same
// argument list -- only the functor carries compiled-in chunk masks instead of run-time ones.
int jit_dense(int op, int pair, const layout_desc& d, const void* a, const void* b, void* out, void* out1, size_t n,
              cudaStream_t s, const launch_geom& g, const pint::cuda::kernels::dense::dense_tuning& tune, bool* handled)
{
    namespace dn = pint::cuda::kernels::dense;
    *handled = false;
    const int fam = pair >= 0 ? FAM_DENSE_PAIR : FAM_DENSE;
    if (!family_applies(d, fam)) return PINT_OK;  // not a whole-chunk layout: the caller's own kernels take it
    jit_module* m = lookup(d, fam);
    if (!m) return PINT_E_JIT;
    if (n == 0) { *handled = true; return PINT_OK; }
    const jit_kernel& k = m->k[pair >= 0 ? pair : op];
    const uint32_t S = sum_bits(d), N = (uint32_t)dense_chunk_words(d);
    if (k.unroll != (int)N || k.block != (int)S || k.vec256 != 8) {  // pint_jit_cfg: {CL::kN, CL::kS, sizeof(mw_geom)}
        std::lock_guard<std::mutex> lock(g_jit_mutex);
        g_last_error = "dense module does not match the layout's chunk geometry";
        return PINT_E_JIT;
    }
    uint32_t T = 0, R = 0;
    dn::choose_tile<uint32_t>(N, tune, &T, &R);
    struct { uint32_t s, chunks_per_thread; } geo = {S, R};  // dense::mw_geom<N>
    uint64_t valid_bits = uint64_t(n) * S;
    size_t total_vec = size_t((valid_bits + 127) / 128);
    const size_t tile_bytes = size_t(T) * R * N * sizeof(uint32_t);
    const size_t ntiles = (total_vec * 16 + tile_bytes - 1) / tile_bytes;
    const size_t smem = 2 * tile_bytes + 16;
    if (ntiles > 0x7FFFFFFFull) return (int)cudaErrorInvalidValue;
    if (smem > 48 * 1024 && g_drv.FuncSetAttribute(k.f128, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, (int)smem) != CUDA_SUCCESS)
        return (int)cudaErrorInvalidValue;
    char empty_fn = 0;
    void* args[8] = {(void*)&a, (void*)&b, (void*)&out, (void*)&out1, &total_vec, &valid_bits, &empty_fn, &geo};
    if (drv_launch(k.f128, unsigned(ntiles), T, (unsigned)smem, s, args, g.pdl != 0) != CUDA_SUCCESS)
        return (int)cudaErrorLaunchFailure;
    count_launch(g);
    *handled = true;
    return PINT_OK;
}

// One operand constant: launches the 16-byte-vector body of [0, n) with the specialised kernel and reports where it
// starts and how many elements it covers; the caller runs its own kernels on what is left on either side.
int jit_scalar_body(int op, bool first, const layout_desc& d, const void* a, uint64_t scalar, void* out, size_t n, cudaStream_t s,
                    const launch_geom& g, size_t* head, size_t* body, bool* handled)
{
    *handled = false;
    jit_module* m = lookup(d, FAM_SCALAR);
    if (!m) return PINT_E_JIT;
    const jit_kernel& k = m->k[(first ? (int)pint::cuda::OP_NUM_BINARY : 0) + op];
    const int wbytes = d.elem_bytes == 8 ? 8 : 4;
    if (k.vec256 != wbytes || k.unroll > 16 || k.block + wbytes > k.unroll) {  // {sizeof(Fn), offsetof(Fn, k), sizeof(word)}
        std::lock_guard<std::mutex> lock(g_jit_mutex);
        g_last_error = "scalar module: unexpected functor layout";
        return PINT_E_JIT;
    }
    const void* ptrs[2] = {a, out};
    size_t h = 0, nvec = 0, tail = 0;
    if (!(split_aligned(d.elem_bytes, n, ptrs, 2, &h, &nvec, &tail, 16) && nvec)) return PINT_OK;
    // the scalar replicated over the processing word, as replicate_scalar does
    uint64_t kw = scalar;
    if (d.elem_bytes == 2) kw = uint32_t(uint16_t(scalar)) * 0x00010001u;
    if (d.elem_bytes == 1) kw = uint32_t(uint8_t(scalar)) * 0x01010101u;
    if (d.elem_bytes == 4) kw = uint32_t(scalar);
    alignas(8) unsigned char fn[16] = {};
    std::memcpy(fn + k.block, &kw, (size_t)wbytes);  // little-endian: the low 4 bytes are the 32-bit word
    const void* in_v = static_cast<const char*>(a) + h * d.elem_bytes;
    void* out_v = static_cast<char*>(out) + h * d.elem_bytes;
    const size_t ntiles = (nvec + 4 * 256 - 1) / (4 * 256);
    void* args[4] = {(void*)&in_v, (void*)&out_v, &nvec, fn};
    if (drv_launch(k.f128, grid_for(ntiles, g), 256, 0, s, args, g.pdl != 0) != CUDA_SUCCESS)
        return (int)cudaErrorLaunchFailure;
    count_launch(g);
    *head = h;
    *body = nvec * (16 / d.elem_bytes);
    *handled = true;
    return PINT_OK;
}

// Interleaved operands: launch_pairs_map's geometry (results 16-byte aligned, pairs 32-byte aligned at the same element)
int jit_pairs_body(int op, const layout_desc& d, const void* pairs, void* out, size_t n, cudaStream_t s, const launch_geom& g,
                   size_t* head, size_t* body, bool* handled)
{
    *handled = false;
    jit_module* m = lookup(d, FAM_PAIRSMAP);
    if (!m) return PINT_E_JIT;
    const jit_kernel& k = m->k[op];
    if (k.unroll != 1) {  // sizeof(Fn): an empty functor
        std::lock_guard<std::mutex> lock(g_jit_mutex);
        g_last_error = "pairsmap module: unexpected functor layout";
        return PINT_E_JIT;
    }
    const size_t eb = d.elem_bytes, per = 16 / eb;
    const uintptr_t off = reinterpret_cast<uintptr_t>(out) & 15u;
    size_t h = off ? (16 - off) / eb : 0;
    if (h > n) h = n;
    const char* pv = static_cast<const char*>(pairs) + 2 * h * eb;
    if (reinterpret_cast<uintptr_t>(pv) & 31u) return PINT_OK;
    size_t nvec = (n - h) / per;
    if (!nvec) return PINT_OK;
    void* ov = static_cast<char*>(out) + h * eb;
    const size_t ntiles = (nvec + 2 * 256 - 1) / (2 * 256);
    char empty_fn = 0;
    void* args[4] = {(void*)&pv, (void*)&ov, &nvec, &empty_fn};
    if (drv_launch(k.f128, grid_for(ntiles, g), 256, 0, s, args, g.pdl != 0) != CUDA_SUCCESS)
        return (int)cudaErrorLaunchFailure;
    count_launch(g);
    *head = h;
    *body = nvec * per;
    *handled = true;
    return PINT_OK;
}

int jit_compile_all(const layout_desc& d)
{
    for (int fam = 0; fam < FAM_COUNT; ++fam)
        if (family_applies(d, fam) && !lookup(d, fam)) return PINT_E_JIT;
    return PINT_OK;
}

// no device needed: fills the disk cache (a build container can warm it for the machines that run)
int jit_prebuild(const layout_desc& d, int major, int minor)
{
    std::lock_guard<std::mutex> lock(g_jit_mutex);
    for (int fam = 0; fam < FAM_COUNT; ++fam) {
        if (!family_applies(d, fam)) continue;
        std::vector<char> cubin;
        std::vector<std::string> lowered;
        if (!build_cubin(d, fam, major, minor, &cubin, &lowered)) return PINT_E_JIT;
    }
    return PINT_OK;
}

bool jit_is_loaded(const layout_desc& d)
{
    std::lock_guard<std::mutex> lock(g_jit_mutex);
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    for (jit_entry* x : g_entries)
        if (x->device == dev && same_layout(x->d, d) && x->fam[FAM_BINARY].state > 0) return true;
    return false;
}

std::string jit_last_error()
{
    std::lock_guard<std::mutex> lock(g_jit_mutex);
    return g_last_error;
}

}  // namespace pint_cuda_impl

extern "C" int pint_cuda_jit_compile(const pint_layout* layout)
{
    pint_cuda_impl::layout_desc d;
    if (int rc = pint_cuda_impl::check_layout(layout, &d)) return rc;
    return pint_cuda_impl::jit_compile_all(d);
}

extern "C" int pint_cuda_jit_prebuild(const pint_layout* layout, int sm_major, int sm_minor)
{
    pint_cuda_impl::layout_desc d;
    if (int rc = pint_cuda_impl::check_layout(layout, &d)) return rc;
    if (sm_major < 8) return PINT_E_ARG;
    return pint_cuda_impl::jit_prebuild(d, sm_major, sm_minor);
}

extern "C" const char* pint_cuda_jit_last_error(void)
{
    static thread_local std::string copy;
    copy = pint_cuda_impl::jit_last_error();
    return copy.c_str();
}
