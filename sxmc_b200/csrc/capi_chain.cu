// capi_chain.cu -- the extern "C" layer of libsxmc_b200.so, part 2: the device-resident MCMC chain
// (MCMC::operator(), src/mcmc.cpp:143-387: tables, step sequencing as CUDA graph / persistent / monotone
// kernels, timed steps), evaluate-only entry points, the NCCL glue and the peer-memory windows of
// sample-sharded fits.
#include "capi_internal.h"

#include <dlfcn.h>

using namespace sxmc;

namespace sxmc { extern float g_last_timed[2]; }

// =========================================================================================
// NCCL glue (dlopen'd so that single-GPU use has no NCCL dependency)
// =========================================================================================
namespace sxmc {
typedef struct { char internal[128]; } nccl_unique_id;
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(nccl_unique_id*) = nullptr;
  int (*CommInitRank)(nccl_comm_t*, int, nccl_unique_id, int) = nullptr;
  int (*CommDestroy)(nccl_comm_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
static NcclApi g_nccl;
static int load_nccl() {
  if (g_nccl.lib) return SXMC_OK;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  void* lib = nullptr;
  for (const char* n : names) { lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (lib) break; }
  if (!lib) { set_error(std::string("cannot load libnccl: ") + dlerror()); return SXMC_ENCCL; }
  g_nccl.GetUniqueId = (int (*)(nccl_unique_id*))dlsym(lib, "ncclGetUniqueId");
  g_nccl.CommInitRank = (int (*)(nccl_comm_t*, int, nccl_unique_id, int))dlsym(lib, "ncclCommInitRank");
  g_nccl.CommDestroy = (int (*)(nccl_comm_t))dlsym(lib, "ncclCommDestroy");
  g_nccl.AllReduce = (int (*)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t))dlsym(lib, "ncclAllReduce");
  g_nccl.GetErrorString = (const char* (*)(int))dlsym(lib, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.CommDestroy) {
    set_error("libnccl lacks a required symbol");
    return SXMC_ENCCL;
  }
  g_nccl.lib = lib;
  return SXMC_OK;
}
static int nccl_fail(int rc, const char* what) {
  set_error(std::string("NCCL error in ") + what + ": " +
            (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?"));
  return SXMC_ENCCL;
}
constexpr int NCCL_UINT32 = 3;   // ncclUint32
constexpr int NCCL_SUM = 0;      // ncclSum
}  // namespace sxmc

namespace sxmc {

static void chain_drop_graph(sxmc_chain* c) {
  if (c->graph_exec) { cudaGraphExecDestroy(c->graph_exec); c->graph_exec = nullptr; }
}

static int chain_build_tables(sxmc_chain* c) {
  const int ns = c->nsignals;
  c->step_bytes = 0;
  for (int b = 0; b < 2; ++b) {
    SXMC_TRY(c->batch[b].reserve(ns));
    unsigned* bins = c->hist_buf[b];
    unsigned* norms = c->hist_buf[b] + (size_t)ns * c->total_nbins;
    unsigned* pbins = c->part_buf[b];
    unsigned* pnorms = c->part_buf[b] + (size_t)ns * c->total_nbins;
    std::vector<NllSignal> sg(ns);
    for (int j = 0; j < ns; ++j) {
      sxmc_hist* h = c->pdfs[j];
      SXMC_TRY(fill_job(h, c->batch[b].host[j], pbins + (size_t)j * c->total_nbins, pnorms + j,
                        c->st_dev->proposed + c->nsources, 1));
      sg[j].bins = bins + (size_t)j * c->total_nbins;
      sg[j].norm = norms + j;
      sg[j].read_bins = h->read_bins_dev;
      sg[j].nexpected = c->nexpected[j];
      sg[j].bin_volume = h->bin_volume;
      sg[j].n_mc = c->n_mc[j];
      sg[j].total_nbins = h->total_nbins;
      sg[j].source_id = c->source_id[j];
    }
    c->batch[b].n_ep = 0;
    plan_launch(c->batch[b], ns, c->tune);
    if (b == 0)
      for (int j = 0; j < ns; ++j)
        c->step_bytes += (long long)c->batch[0].host[j].nwtiles * WT_SAMPLES * 4ll *
                         std::max(1, c->batch[0].host[j].nslots);
    SXMC_TRY(upload_batch(c->batch[b], ns, c->stream));
    SXMC_CUDA(cudaMemcpyAsync(c->sigs_dev[b], sg.data(), sizeof(NllSignal) * ns, cudaMemcpyHostToDevice, c->stream));
  }
  SXMC_CUDA(cudaStreamSynchronize(c->stream));    // `sg` and the pinned staging tables may be reused after this
  return SXMC_OK;
}

// in_step: the launch follows the previous step's lookup kernel in the stream -- its prologue may
// overlap that kernel's tail (programmatic dependent launch)
static void chain_p2p_params(sxmc_chain* c, int b, P2PParams& W) {
  memset(&W, 0, sizeof W);
  for (int p = 0; p < c->p2p_nranks; ++p) {
    W.part[p] = (const unsigned*)c->p2p_peer[p];
    W.flags[p] = (unsigned*)c->p2p_peer[p] + 4 * c->p2p_pitch;
  }
  W.sum = c->hist_buf[b];
  W.sum_offset = (2 + (size_t)b) * c->p2p_pitch;
  // form 3 (exchange fused into a rank-sharded lookup) applies to steps only: every other
  // evaluation, and the last step of a call, exchanges whole histograms in the automatic form
  W.twoshot = (c->tune.p2p_form == 1 || c->tune.p2p_form == 2) ? (c->tune.p2p_form == 2) : (c->p2p_nranks > 2);
  W.seq = c->p2p_state; W.ticket = c->p2p_state + 1; W.abort_flag = c->p2p_state + 2;
  W.words = c->hist_words; W.pitch = c->p2p_pitch; W.parity = b;
  W.peer_timeout_ns = c->p2p_timeout_ns;
  W.nranks = c->p2p_nranks; W.rank = c->p2p_rank;
}

// A sharded step may run its exchange inside the lookup launch (nll_sharded_kernels.cu) when the
// unsharded chain would take the grouped lookup over many bins -- the launch whose partial sums
// the fused kernel reproduces block for block.
static bool chain_fused_ok(const sxmc_chain* c) {
  if (c->p2p_nranks < 2 || !c->mult_dev || c->nsignals > 128) return false;
  const int form = c->tune.p2p_form;
  if (form == 1 || form == 2) return false;             // a whole-histogram form was asked for
  if (c->total_nbins <= 4096 && c->hist_words <= 65536) return false;   // launch_nll: one-block lookup
  const int g = nll_sharded_logical_grid(c->total_nbins);
  return g >= c->p2p_nranks && g <= P2P_PARTIALS_MAX;
}

// fused: the exchange happens inside the lookup launch that follows (chain_nll)
static int chain_bin(sxmc_chain* c, int b, cudaEvent_t after_bin = nullptr, bool in_step = false, bool fused = false) {
  BinLaunch BL = c->batch[b].launch;
  BL.pdl = in_step && !c->tune.no_pdl;
  BL.cta_split = c->tune.cta_split;
  SXMC_TRY(launch_bin(BL, c->stream));
  c->launches += 1;
  if (after_bin) SXMC_CUDA(cudaEventRecord(after_bin, c->stream));
  if (fused) return SXMC_OK;
  if (c->p2p_nranks > 1) {
    P2PParams W;
    chain_p2p_params(c, b, W);
    SXMC_TRY(launch_hist_allreduce_p2p(W, c->stream));
    c->launches += 1;
  } else if (c->comm && c->comm->nranks > 1) {
    int rc = g_nccl.AllReduce(c->hist_buf[b], c->hist_buf[b], c->hist_words, NCCL_UINT32, NCCL_SUM,
                              c->comm->comm, c->stream);
    if (rc) return nccl_fail(rc, "ncclAllReduce");
    c->launches += 1;
  }
  return SXMC_OK;
}

static int chain_nll(sxmc_chain* c, int b, int mode, bool zero_other, float* lut, bool fused = false) {
  if (fused) {
    ShardNllParams Q;
    memset(&Q, 0, sizeof Q);
    chain_p2p_params(c, b, Q.W);
    Q.sigs = c->sigs_dev[b];
    Q.st = c->st_dev;
    Q.mult = c->mult_dev;
    Q.zero_ptr = c->part_buf[1 - b];
    Q.zero_words = c->hist_words;
    Q.partials_offset = 4 * c->p2p_pitch + 2 * P2P_MAX_RANKS;
    Q.ns = c->nsignals; Q.np = c->nparameters; Q.nsources = c->nsources; Q.nbins = c->total_nbins; Q.mode = mode;
    SXMC_TRY(launch_nll_sharded(Q, !c->tune.no_pdl, c->stream));
    c->launches += 1;
    return SXMC_OK;
  }
  NllLaunch L;
  L.sigs_dev = c->sigs_dev[b];
  L.norms_dev = c->hist_buf[b] + (size_t)c->nsignals * c->total_nbins;
  L.nsignals = c->nsignals;
  L.nparameters = c->nparameters;
  L.nsources = c->nsources;
  L.nevents = (long long)c->nevents;
  L.state = c->st_dev;
  L.partials = c->partials;
  L.ticket = c->ticket;
  L.mode = mode;
  L.lut = lut;
  L.zero_ptr = zero_other ? c->part_buf[1 - b] : nullptr;
  L.zero_words = zero_other ? c->hist_words : 0;
  L.grid = c->nll_grid;
  L.mult = c->mult_dev;
  L.mult_nbins = c->total_nbins;
  L.ptable = nullptr;
  // a step that re-binned: the lookup's prologue overlaps the binning kernel's tail; the early clear
  // of the other histogram buffer only when no peer can still be reading it
  L.pdl = mode == NLL_MODE_STEP && zero_other && !c->tune.no_pdl;
  L.zero_early = L.pdl && !((c->comm && c->comm->nranks > 1) || c->p2p_nranks > 1);
  // events looked up one by one (signals of several datasets): eval_pdf once per (signal, bin) first
  if (c->ptable_dev && !c->mult_dev && !lut && mode == NLL_MODE_STEP && c->nevents * (size_t)c->nsignals >= 4096) {
    if (zero_other || !c->ptable_valid) {        // zero_other == the step re-binned
      SXMC_TRY(launch_pdf_table(L.sigs_dev, L.norms_dev, c->nsignals, c->total_nbins, c->ptable_dev, c->stream));
      c->launches += 1;
      c->ptable_valid = !zero_other;
    }
    L.ptable = c->ptable_dev;
  }
  SXMC_TRY(launch_nll(L, c->stream));
  c->launches += 1;
  return SXMC_OK;
}

}  // namespace sxmc

extern "C" {

int sxmc_hists_eval_host(sxmc_hist* const* hs, int n, const double* params, int nparams,
                         long param_offset, unsigned* norms_out, float* pdf_out) {
  if (n <= 0) return SXMC_OK;
  SXMC_TRY(ensure_init());
  cudaStream_t stream = hs[0]->stream;
  double* p_dev = nullptr;
  unsigned* n_dev = nullptr;
  float* pdf_dev = nullptr;
  size_t npts = hs[0]->npoints;
  for (int j = 1; j < n; ++j) npts = std::max(npts, hs[j]->npoints);
  cudaError_t e = cudaMalloc(&p_dev, sizeof(double) * std::max(1, nparams));
  if (e == cudaSuccess) e = cudaMalloc(&n_dev, sizeof(unsigned) * n);
  if (e == cudaSuccess && pdf_out && npts) e = cudaMalloc(&pdf_dev, sizeof(float) * npts * n);
  int rc = SXMC_OK;
  if (e == cudaSuccess && nparams > 0)
    e = cudaMemcpyAsync(p_dev, params, sizeof(double) * nparams, cudaMemcpyHostToDevice, stream);
  if (e == cudaSuccess) {
    for (int j = 0; j < n; ++j) {
      sxmc_hist_set_param_buffer(hs[j], p_dev, param_offset, 1);
      sxmc_hist_set_norm_buffer(hs[j], n_dev, j);
      sxmc_hist_set_pdf_buffer(hs[j], pdf_dev, (long)(j * npts), 1);
    }
    rc = sxmc_hists_eval_async(hs, n, pdf_dev != nullptr);
    if (!rc) e = cudaMemcpyAsync(norms_out, n_dev, sizeof(unsigned) * n, cudaMemcpyDeviceToHost, stream);
    if (!rc && e == cudaSuccess && pdf_dev)
      e = cudaMemcpyAsync(pdf_out, pdf_dev, sizeof(float) * npts * n, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  }
  for (int j = 0; j < n; ++j) {   // the temporaries die here: unbind them
    sxmc_hist_set_param_buffer(hs[j], nullptr, 0, 1);
    sxmc_hist_set_norm_buffer(hs[j], nullptr, 0);
    sxmc_hist_set_pdf_buffer(hs[j], nullptr, 0, 1);
  }
  cudaFree(p_dev); cudaFree(n_dev); cudaFree(pdf_dev);
  if (e != cudaSuccess) return cuda_fail(e, "sxmc_hists_eval_host", __FILE__, __LINE__);
  return rc;
}

int sxmc_nll_eval(const float* lut_dev, size_t nevents, size_t nsignals, const double* pars_dev,
                  size_t nparameters, size_t nsources, const double* means_dev,
                  const double* sigmas_dev, const double* nexpected_dev, const unsigned* n_mc_dev,
                  const short* source_id_dev, const unsigned* norms_dev, double* nll_out_host,
                  double* event_sum_out_host) {
  SXMC_TRY(ensure_init());
  double* out2 = nullptr;
  SXMC_CUDA(cudaMalloc(&out2, 2 * sizeof(double)));
  int rc = launch_nll_lut(lut_dev, nevents, nsignals, pars_dev, nparameters, nsources, means_dev,
                          sigmas_dev, nexpected_dev, n_mc_dev, source_id_dev, norms_dev, out2, 0);
  double h[2] = {0, 0};
  cudaError_t e = cudaMemcpy(h, out2, sizeof h, cudaMemcpyDeviceToHost);
  cudaFree(out2);
  if (e != cudaSuccess) return cuda_fail(e, "sxmc_nll_eval", __FILE__, __LINE__);
  if (nll_out_host) *nll_out_host = h[0];
  if (event_sum_out_host) *event_sum_out_host = h[1];
  return rc;
}

// ---------------------------------------------------------------------------------------
// MCMC
// ---------------------------------------------------------------------------------------
int sxmc_chain_create(const sxmc_chain_config* cfg, sxmc_chain** out) {
  if (!cfg || !out) SXMC_INVAL("sxmc_chain_create: NULL argument");
  *out = nullptr;
  if (cfg->nparameters <= 0 || cfg->nparameters > SXMC_MAX_PARAMETERS)
    SXMC_INVAL("sxmc_chain_create: nparameters out of range (SXMC_MAX_PARAMETERS)");
  if (cfg->nsignals <= 0 || cfg->nsignals > 128) SXMC_INVAL("sxmc_chain_create: nsignals out of range (1..128)");
  if (cfg->nsources <= 0 || cfg->nsources > cfg->nparameters) SXMC_INVAL("sxmc_chain_create: nsources out of range");
  SXMC_TRY(ensure_init());
  sxmc_chain* c = new sxmc_chain;
  c->tune = g_tune;
  c->nsources = cfg->nsources; c->nsignals = cfg->nsignals; c->nparameters = cfg->nparameters;
  c->nobs = cfg->nobservables; c->nsystematics = cfg->nsystematics;
  c->systematics_fixed = cfg->systematics_fixed != 0;
  c->means.assign(cfg->parameter_means, cfg->parameter_means + cfg->nparameters);
  c->sigmas.assign(cfg->parameter_sigma, cfg->parameter_sigma + cfg->nparameters);
  c->fixed.assign(cfg->parameter_fixed, cfg->parameter_fixed + cfg->nparameters);
  c->pdfs.assign(cfg->pdfs, cfg->pdfs + cfg->nsignals);
  c->nexpected.assign(cfg->nexpected, cfg->nexpected + cfg->nsignals);
  c->n_mc.assign(cfg->n_mc, cfg->n_mc + cfg->nsignals);
  c->source_id.assign(cfg->source_id, cfg->source_id + cfg->nsignals);
  c->total_nbins = c->pdfs[0]->total_nbins;
  for (int j = 0; j < c->nsignals; ++j) {
    if (c->pdfs[j]->total_nbins != c->total_nbins || c->pdfs[j]->nobs != c->nobs) {
      delete c;
      SXMC_INVAL("sxmc_chain_create: all signal PDFs must share the observables' binning");
    }
    if (c->source_id[j] < 0 || c->source_id[j] >= c->nsources) {
      delete c;
      SXMC_INVAL("sxmc_chain_create: source_id out of range");
    }
  }
  c->hist_words = (size_t)c->nsignals * ((size_t)c->total_nbins + 1);
  cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaMalloc(&c->st_dev, sizeof(ChainState));
  if (e == cudaSuccess) e = cudaMallocHost(&c->st_host, sizeof(ChainState));
  for (int b = 0; b < 2 && e == cudaSuccess; ++b) {
    const size_t padded = (c->hist_words + 31) & ~(size_t)31;     // whole 128-byte lines (staged peer exchange copies uint4)
    e = cudaMalloc(&c->hist_buf[b], sizeof(unsigned) * padded);
    if (e == cudaSuccess) e = cudaMemset(c->hist_buf[b], 0, sizeof(unsigned) * padded);
    c->part_buf[b] = c->hist_buf[b];
    c->hist_own[b] = c->hist_buf[b];
    if (e == cudaSuccess) e = cudaMalloc(&c->sigs_dev[b], sizeof(NllSignal) * c->nsignals);
  }
  if (e == cudaSuccess) e = cudaMalloc(&c->partials, sizeof(double) * 148 * 8);
  if (e == cudaSuccess) e = cudaMalloc(&c->ticket, sizeof(unsigned));
  if (e == cudaSuccess) e = cudaMalloc(&c->bcast_dev, sizeof(double) * 2 * SXMC_MAX_PARAMETERS);
  if (e == cudaSuccess) e = cudaMalloc(&c->sync_dev, sizeof(unsigned) * 4);
  if (e == cudaSuccess) e = cudaMallocHost(&c->sync_host, sizeof(unsigned) * 4);
  if (e == cudaSuccess) e = cudaMemset(c->ticket, 0, sizeof(unsigned));
  if (e == cudaSuccess) e = cudaEventCreate(&c->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&c->ev1);
  if (e != cudaSuccess) { sxmc_chain_destroy(c); return cuda_fail(e, "sxmc_chain_create", __FILE__, __LINE__); }
  *out = c;
  return SXMC_OK;
}

int sxmc_chain_destroy(sxmc_chain* c) {
  if (!c) return SXMC_OK;
  if (c->stream) cudaStreamSynchronize(c->stream);
  chain_drop_graph(c);
  cudaFree(c->flush_buf);
  cudaFree(c->bcast_dev); cudaFree(c->sync_dev); cudaFree(c->stamps_dev);
  cudaFree(c->mult_dev); cudaFree(c->nevents_dev); cudaFree(c->data_dev);
  for (int p = 0; p < c->p2p_nranks; ++p)
    if (p != c->p2p_rank && c->p2p_peer[p]) cudaIpcCloseMemHandle(c->p2p_peer[p]);
  cudaFree(c->p2p_window); cudaFree(c->p2p_state);
  if (c->p2p_state_host) cudaFreeHost(c->p2p_state_host);
  delete c->mono_plan;
  if (c->sync_host) cudaFreeHost(c->sync_host);
  for (auto& e : c->tev) cudaEventDestroy(e);
  cudaFree(c->st_dev);
  if (c->st_host) cudaFreeHost(c->st_host);
  for (int b = 0; b < 2; ++b) { cudaFree(c->hist_own[b]); cudaFree(c->sigs_dev[b]); c->batch[b].release(); }
  cudaFree(c->partials); cudaFree(c->ticket); cudaFree(c->lut_dev); cudaFree(c->rows_dev); cudaFree(c->ptable_dev);
  cudaFree(c->nll_trace_dev); cudaFree(c->prop_trace_dev); cudaFree(c->tape_n_dev); cudaFree(c->tape_u_dev);
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
  return SXMC_OK;
}

int sxmc_chain_set_seed(sxmc_chain* c, uint64_t seed, uint64_t stream_id) {
  c->seed = seed; c->stream_id = stream_id; return SXMC_OK;
}

int sxmc_chain_set_tapes(sxmc_chain* c, const double* normals, size_t n_normals,
                         const double* uniforms, size_t n_uniforms) {
  cudaFree(c->tape_n_dev); cudaFree(c->tape_u_dev);
  c->tape_n_dev = nullptr; c->tape_u_dev = nullptr; c->n_tape_n = 0; c->n_tape_u = 0;
  if (normals && uniforms) {
    SXMC_CUDA(cudaMalloc(&c->tape_n_dev, sizeof(double) * std::max<size_t>(1, n_normals)));
    SXMC_CUDA(cudaMalloc(&c->tape_u_dev, sizeof(double) * std::max<size_t>(1, n_uniforms)));
    SXMC_CUDA(cudaMemcpy(c->tape_n_dev, normals, sizeof(double) * n_normals, cudaMemcpyHostToDevice));
    SXMC_CUDA(cudaMemcpy(c->tape_u_dev, uniforms, sizeof(double) * n_uniforms, cudaMemcpyHostToDevice));
    c->n_tape_n = n_normals; c->n_tape_u = n_uniforms;
  }
  return SXMC_OK;
}

int sxmc_chain_set_comm(sxmc_chain* c, sxmc_comm* comm) { chain_drop_graph(c); c->comm = comm; return SXMC_OK; }
int sxmc_chain_set_mode(sxmc_chain* c, int mode) { c->mode = mode; return SXMC_OK; }
int sxmc_chain_phase_stamps(sxmc_chain* c, int enable, unsigned long long* out64x8) {
  if (enable && !c->stamps_dev) {
    SXMC_CUDA(cudaMalloc(&c->stamps_dev, sizeof(unsigned long long) * 2048));
    SXMC_CUDA(cudaMemset(c->stamps_dev, 0, sizeof(unsigned long long) * 2048));
  }
  if (out64x8 && c->stamps_dev)
    SXMC_CUDA(cudaMemcpy(out64x8, c->stamps_dev, sizeof(unsigned long long) * 2048, cudaMemcpyDeviceToHost));
  if (!enable && c->stamps_dev) { cudaFree(c->stamps_dev); c->stamps_dev = nullptr; }
  return SXMC_OK;
}
int sxmc_chain_enable_trace(sxmc_chain* c, int on) { c->trace = on != 0; return SXMC_OK; }

int sxmc_chain_set_data(sxmc_chain* c, const float* data, size_t data_nfloats) {
  if (data_nfloats % (size_t)(c->nobs + 1) != 0)
    SXMC_INVAL("Number of entries in evaluation points array not divisible by number of observables.");
  c->nevents = data_nfloats / (size_t)(c->nobs + 1);
  // One upload of the data events for all PDFs, then p->SetEvalPoints(data) for every pdf
  // (src/mcmc.cpp:235) as launches on the chain's stream.  Buffers only ever grow, so that a fit
  // re-run on new data of the same size keeps its tables and its captured step graph.
  if (data_nfloats > c->data_cap || !c->data_dev) {
    cudaFree(c->data_dev);
    c->data_dev = nullptr;
    c->data_cap = 0;
    SXMC_CUDA(cudaMalloc(&c->data_dev, sizeof(float) * std::max<size_t>(1, data_nfloats)));
    c->data_cap = std::max<size_t>(1, data_nfloats);
  }
  if (data_nfloats)
    SXMC_CUDA(cudaMemcpyAsync(c->data_dev, data, sizeof(float) * data_nfloats, cudaMemcpyHostToDevice, c->stream));
  for (int j = 0; j < c->nsignals; ++j) SXMC_TRY(hist_set_eval_points_dev(c->pdfs[j], c->data_dev, c->nevents, c->stream));
  c->nll_grid = nll_grid_for((long long)c->nevents);
  const size_t lut_need = std::max<size_t>(1, c->nevents * c->nsignals);
  if (lut_need > c->lut_cap || !c->lut_dev) {
    cudaFree(c->lut_dev);
    c->lut_dev = nullptr;
    c->lut_cap = 0;
    SXMC_CUDA(cudaMalloc(&c->lut_dev, sizeof(float) * lut_need));
    c->lut_cap = lut_need;
  }
  // grouped lookup: events per bin, when all signals read one dataset
  bool shared = true;
  for (int j = 1; j < c->nsignals; ++j) shared = shared && c->pdfs[j]->dataset == c->pdfs[0]->dataset;
  c->mult_fits_table = (size_t)c->nsignals * c->total_nbins <= (size_t)BATCH_PDF_TABLE;
  if (shared && !c->tune.no_group) {
    if (!c->mult_dev) SXMC_CUDA(cudaMalloc(&c->mult_dev, sizeof(unsigned) * (size_t)c->total_nbins));
    SXMC_TRY(launch_event_mult(c->pdfs[0]->read_bins_dev, (long long)c->nevents, c->total_nbins, c->mult_dev, c->stream));
  } else {
    if (c->mult_dev) { cudaFree(c->mult_dev); c->mult_dev = nullptr; }
    if (!c->ptable_dev)     // per-event lookup: the step's PDF values per (signal, bin)
      SXMC_CUDA(cudaMalloc(&c->ptable_dev, sizeof(double) * (size_t)c->nsignals * (size_t)c->total_nbins));
  }
  if (!c->nevents_dev) SXMC_CUDA(cudaMalloc(&c->nevents_dev, sizeof(long long)));
  const long long ne = (long long)c->nevents;
  SXMC_CUDA(cudaMemcpyAsync(c->nevents_dev, &ne, sizeof ne, cudaMemcpyHostToDevice, c->stream));
  // everything the uploaded tables and the captured graph depend on
  std::vector<const void*> sig, lsig;
  sig.push_back((const void*)(uintptr_t)c->nevents);
  sig.push_back(c->lut_dev); sig.push_back(c->mult_dev);
  sig.push_back(c->hist_buf[0]); sig.push_back(c->hist_buf[1]); sig.push_back(c->part_buf[0]); sig.push_back(c->part_buf[1]);
  for (int j = 0; j < c->nsignals; ++j) {
    sxmc_hist* h = c->pdfs[j];
    SXMC_TRY(ensure_prepared(h));
    sig.push_back(h->read_bins_dev);
    lsig.push_back(h->cols_dev);
    lsig.push_back((const void*)(uintptr_t)(h->systs.size() * 16 + (size_t)h->primary + 1));
  }
  sig.insert(sig.end(), lsig.begin(), lsig.end());
  if (sig != c->table_sig) {
    chain_drop_graph(c);   // the captured launches bake nevents and the table pointers
    SXMC_TRY(chain_build_tables(c));
    c->table_sig = sig;
  }
  if (lsig != c->layout_sig) {
    c->layout_sig = lsig;
    c->mono_checked = false;
    delete c->mono_plan;
    c->mono_plan = nullptr;
  }
  SXMC_CUDA(cudaStreamSynchronize(c->stream));   // `data` may be freed by the caller after this returns
  c->have_data = true;
  return SXMC_OK;
}

}  // extern "C"

namespace sxmc {

// Host-side initial state: the block of src/mcmc.cpp:152-228.
void init_state_common(ChainState& s, int np, int nsources, int nsignals, const double* means,
                              const double* sigmas, const int* fixed, unsigned nsteps, float burnin_fraction,
                              int debug_mode) {
  memset(&s, 0, sizeof s);
  int nfloat = 0;
  for (int i = 0; i < np; ++i) {
    s.current[i] = means[i];
    s.proposed[i] = means[i];
    s.means[i] = means[i];
    s.sigmas[i] = sigmas[i];
    s.fixed[i] = fixed[i];
    nfloat += fixed[i] ? 0 : 1;
  }
  const float scale_factor = 2.4 * 2.4 / nfloat;   // Haario, 2001 (src/mcmc.cpp:199)
  s.scale_factor = scale_factor;
  for (int i = 0; i < np; ++i) {
    if (fixed[i]) { s.jump_width[i] = -1; continue; }
    float mean = means[i];
    float sigma = sigmas[i];
    float width = 0.1;
    if (sigma > 0) {
      width = sigma;
    } else if (i < nsignals) {     // sic: nsignals, not nsources (src/mcmc.cpp:217)
      float m = std::max(mean, (float)10);
      width = sqrt(m) / m;
    } else {
      width = sqrt(std::max(mean, (float)1));
    }
    s.jump_width[i] = 0.1 * width * scale_factor;
  }
  s.nparameters = np;
  s.nsources = nsources;
  s.nsignals = nsignals;
  s.debug_mode = debug_mode;
  s.step = 0;
  s.burnin_steps = nsteps * burnin_fraction;        // src/mcmc.cpp:152
}

static void chain_init_state(sxmc_chain* c, ChainState& s, unsigned nsteps, float burnin_fraction,
                             int debug_mode) {
  init_state_common(s, c->nparameters, c->nsources, c->nsignals, c->means.data(), c->sigmas.data(),
                    c->fixed.data(), nsteps, burnin_fraction, debug_mode);
  s.seed = c->seed;
  s.stream_id = c->stream_id;
  s.tape_normals = c->tape_n_dev; s.n_normals = c->n_tape_n;
  s.tape_uniforms = c->tape_u_dev; s.n_uniforms = c->n_tape_u;
  s.rows = c->rows_dev;
  s.row_cap = c->row_cap;
  s.nll_trace = c->trace ? c->nll_trace_dev : nullptr;
  s.prop_trace = c->trace ? c->prop_trace_dev : nullptr;
}

static int chain_push_state(sxmc_chain* c) {
  SXMC_CUDA(cudaMemcpyAsync(c->st_dev, c->st_host, sizeof(ChainState), cudaMemcpyHostToDevice, c->stream));
  return SXMC_OK;
}
static int chain_pull_state(sxmc_chain* c) {
  SXMC_CUDA(cudaMemcpyAsync(c->st_host, c->st_dev, sizeof(ChainState), cudaMemcpyDeviceToHost, c->stream));
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  return SXMC_OK;
}

// after a synchronisation: did a peer fail to show up in a peer-memory exchange?
static int chain_p2p_check(sxmc_chain* c) {
  if (c->p2p_nranks <= 1) return SXMC_OK;
  SXMC_CUDA(cudaMemcpy(c->p2p_state_host, c->p2p_state, sizeof(unsigned) * 4, cudaMemcpyDeviceToHost));
  if (c->p2p_state_host[2]) {
    set_error("peer-memory histogram exchange timed out waiting for another rank");
    return SXMC_ESTATE;
  }
  return SXMC_OK;
}

// Fixed systematics: every step reads the initial evaluation (buffer 0), so the PDF-value table of
// the per-event lookup is computed once, outside any graph capture.
static int chain_fixed_table(sxmc_chain* c, bool rebin) {
  if (rebin || c->mult_dev || !c->ptable_dev || c->ptable_valid) return SXMC_OK;
  if (c->nevents * (size_t)c->nsignals < 4096) return SXMC_OK;
  SXMC_TRY(launch_pdf_table(c->sigs_dev[0], c->hist_buf[0] + (size_t)c->nsignals * c->total_nbins, c->nsignals,
                            c->total_nbins, c->ptable_dev, c->stream));
  c->launches += 1;
  c->ptable_valid = true;
  return SXMC_OK;
}

// one MCMC step enqueued on the chain's stream: [bin -> (allreduce)] -> lookup+nll+decide+propose
static int chain_enqueue_step(sxmc_chain* c, bool rebin, int b, bool fused = false) {
  fused = fused && rebin;
  if (rebin) SXMC_TRY(chain_bin(c, b, nullptr, true, fused));
  SXMC_TRY(chain_nll(c, b, NLL_MODE_STEP, rebin, nullptr, fused));
  return SXMC_OK;
}

}  // namespace sxmc

extern "C" {

int sxmc_chain_begin(sxmc_chain* c, const float* data, size_t data_nfloats, unsigned nsteps,
                     float burnin_fraction, int debug_mode) {
  if (!c) SXMC_INVAL("sxmc_chain_begin: NULL chain");
  SXMC_TRY(sxmc_chain_set_data(c, data, data_nfloats));
  const int np = c->nparameters;
  // chain buffer: every step fits (the reference drains a sync_interval-sized buffer instead)
  if (c->row_cap < (long long)nsteps || !c->rows_dev) {
    cudaFree(c->rows_dev);
    c->rows_dev = nullptr;
    c->row_cap = std::max<long long>(1, nsteps);
    SXMC_CUDA(cudaMalloc(&c->rows_dev, sizeof(float) * (size_t)c->row_cap * (np + 1)));
  }
  if (c->trace && c->trace_steps < nsteps) {
    cudaFree(c->nll_trace_dev); cudaFree(c->prop_trace_dev);
    c->nll_trace_dev = nullptr; c->prop_trace_dev = nullptr;
    SXMC_CUDA(cudaMalloc(&c->nll_trace_dev, sizeof(double) * std::max(1u, nsteps)));
    SXMC_CUDA(cudaMalloc(&c->prop_trace_dev, sizeof(double) * std::max(1u, nsteps) * np));
    c->trace_steps = nsteps;
  }
  chain_init_state(c, *c->st_host, nsteps, burnin_fraction, debug_mode);
  SXMC_TRY(chain_push_state(c));
  if (c->p2p_state)     // a timed-out exchange of an earlier run must not poison this one
    SXMC_CUDA(cudaMemsetAsync(c->p2p_state + 2, 0, sizeof(unsigned), c->stream));
  c->launches = 0;
  c->loop_ms = 0;
  c->steps_done = 0;
  c->steps_planned = nsteps;

  // initial evaluation at the means (src/mcmc.cpp:233-252): always bins, whatever is fixed
  for (int b = 0; b < 2; ++b) {
    SXMC_CUDA(cudaMemsetAsync(c->hist_buf[b], 0, sizeof(unsigned) * c->hist_words, c->stream));
    if (c->part_buf[b] != c->hist_buf[b])
      SXMC_CUDA(cudaMemsetAsync(c->part_buf[b], 0, sizeof(unsigned) * c->hist_words, c->stream));
  }
  SXMC_TRY(chain_bin(c, 0));
  SXMC_TRY(chain_nll(c, 0, NLL_MODE_INIT, false, nullptr));
  c->ptable_valid = false;
  c->last_buf = 0;
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  return chain_p2p_check(c);
}

// Advance the chain by nsteps (the body of the loop at src/mcmc.cpp:261-378), device-resident.
int sxmc_chain_advance(sxmc_chain* c, unsigned nsteps) {
  if (!c || !c->have_data) { set_error("sxmc_chain_advance: sxmc_chain_begin was not called"); return SXMC_ESTATE; }
  if ((unsigned long long)c->steps_done + nsteps > c->steps_planned) {
    set_error("sxmc_chain_advance: more steps than sxmc_chain_begin announced");
    return SXMC_ESTATE;
  }
  const bool rebin = c->nsystematics > 0 && !c->systematics_fixed;   // src/mcmc.cpp:264
  const bool has_comm = (c->comm && c->comm->nranks > 1) || c->p2p_nranks > 1;
  SXMC_TRY(chain_fixed_table(c, rebin));
  const unsigned GRAPH_STEPS = 16;   // even: the histogram double-buffer parity is preserved
  unsigned left = nsteps;
  // Small fits: the whole segment in one cooperative launch (chain_kernels.cu).  Chosen when a
  // step is too short for per-kernel launches to be hidden (the samples fit in L2).
  c->used_persistent = false;
  const BinLaunch& BL = c->batch[0].launch;
  const bool persistent_ok = BL.fast && BL.njobs <= FAST_MAX_JOBS && !has_comm && c->nsignals <= 128 &&
                             c->pdfs.size() && nsteps > 0;
  // Monotone fits (one observable, shift / scale / ctscale only): the edge-search kernel of
  // batch_mono_kernels.cu with a batch of one chain.
  if ((c->mode == 0 || c->mode == 4) && !has_comm && nsteps > 0 && c->pdfs.size()) {
    if (!c->mono_checked) {
      c->mono_checked = true;
      bool mono = BL.fast && BL.nobs == 1 && BL.nslots == 1;
      for (int j = 0; j < c->nsignals && mono; ++j) {
        sxmc_hist* h = c->pdfs[j];
        if (!h->sort || !h->prepared || h->primary != 0) mono = false;
        for (const SystDesc& d : h->systs)
          if (d.type == SXMC_SYST_RESOLUTION_SCALE || d.obs != 0 || d.npars > 1) mono = false;
        bool bad = false;
        if (mono) SXMC_TRY(column_has_nan_or_neginf(h->cols_dev, h->col_pitch, &bad, c->stream));
        if (bad) mono = false;
      }
      if (mono) {
        std::vector<unsigned long long> nsamp(c->nsignals);
        for (int j = 0; j < c->nsignals; ++j) nsamp[j] = c->pdfs[j]->nsamples;
        c->mono_plan = new MonoPlan;
        if (plan_chain_batch_mono(BL, nsamp.data(), 1, c->mono_plan) != SXMC_OK) { delete c->mono_plan; c->mono_plan = nullptr; }
      }
    }
    if (c->mono_plan && !c->tune.no_mono) {
      BatchParams Q;
      memset(&Q, 0, sizeof Q);
      Q.st = c->st_dev; Q.sigs = c->sigs_dev[0]; Q.hist = c->hist_buf[0]; Q.hist_words = c->hist_words;
      Q.nevents = c->nevents_dev; Q.sync = c->sync_dev;
      Q.nchains = 1; Q.ns = c->nsignals; Q.np = c->nparameters; Q.nsources = c->nsources;
      Q.nsteps = nsteps; Q.do_init = 0; Q.rebin = rebin ? 1 : 0;
      Q.mult = c->mult_fits_table ? c->mult_dev : nullptr; Q.mono = 1;
      if (rebin) SXMC_CUDA(cudaMemsetAsync(c->hist_buf[0], 0, sizeof(unsigned) * c->hist_words, c->stream));
      else if (c->last_buf != 0) {   // fixed systematics: the lookups read the initial evaluation
        SXMC_CUDA(cudaMemcpyAsync(c->hist_buf[0], c->hist_buf[c->last_buf], sizeof(unsigned) * c->hist_words,
                                  cudaMemcpyDeviceToDevice, c->stream));
      }
      SXMC_CUDA(cudaMemsetAsync(c->sync_dev, 0, sizeof(unsigned) * 4, c->stream));
      SXMC_CUDA(cudaEventRecord(c->ev0, c->stream));
      SXMC_TRY(launch_chain_batch_mono(*c->mono_plan, Q, c->stream));
      SXMC_CUDA(cudaEventRecord(c->ev1, c->stream));
      SXMC_CUDA(cudaMemcpyAsync(c->sync_host, c->sync_dev, sizeof(unsigned) * 4, cudaMemcpyDeviceToHost, c->stream));
      SXMC_CUDA(cudaStreamSynchronize(c->stream));
      if (c->sync_host[3]) { set_error("monotone chain kernel timed out at a grid barrier"); return SXMC_ESTATE; }
      float ms = 0;
      SXMC_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
      c->loop_ms += ms;
      c->launches += 1;
      c->steps_done += nsteps;
      c->last_buf = 0;
      c->last_path = 4;
      return SXMC_OK;
    }
    if (c->mode == 4) { set_error("sxmc_chain_set_mode(4): the fit is not monotone (or does not fit on chip)"); return SXMC_ESTATE; }
  }
  const bool want_persistent = c->mode == 2 || (c->mode == 0 && c->step_bytes <= c->tune.persistent_max_bytes);
  if (persistent_ok && want_persistent) {
    PersistParams Q;
    memset(&Q, 0, sizeof Q);
    Q.st = c->st_dev;
    Q.sigs = c->sigs_dev[0];
    Q.hist_buf[0] = c->hist_buf[0];
    Q.hist_buf[1] = c->hist_buf[1];
    Q.hist_words = c->hist_words;
    Q.partials = c->partials;
    Q.bcast = c->bcast_dev;
    Q.sync = c->sync_dev;
    Q.nevents = (long long)c->nevents;
    Q.ns = c->nsignals; Q.np = c->nparameters; Q.nsources = c->nsources;
    Q.nsteps = nsteps;
    Q.step0 = c->steps_done;
    Q.rebin = rebin ? 1 : 0;
    Q.stamps = c->stamps_dev;
    SXMC_CUDA(cudaMemsetAsync(c->sync_dev, 0, sizeof(unsigned) * 4, c->stream));
    SXMC_CUDA(cudaEventRecord(c->ev0, c->stream));
    int rc = launch_chain_persistent(BL, Q, c->stream);
    if (rc == SXMC_OK) {
      SXMC_CUDA(cudaEventRecord(c->ev1, c->stream));
      SXMC_CUDA(cudaMemcpyAsync(c->sync_host, c->sync_dev, sizeof(unsigned) * 4, cudaMemcpyDeviceToHost, c->stream));
      SXMC_CUDA(cudaStreamSynchronize(c->stream));
      if (c->sync_host[3]) { set_error("persistent chain kernel timed out at a grid barrier"); return SXMC_ESTATE; }
      float ms = 0;
      SXMC_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
      c->loop_ms += ms;
      c->launches += 1;
      c->steps_done += nsteps;
      if (rebin && c->steps_done > 0) c->last_buf = (int)(c->steps_done & 1);
      c->used_persistent = true;
      c->last_path = 2;
      return SXMC_OK;
    }
    if (c->mode == 2) return rc;     // explicitly requested: report why it cannot run
    cudaGetLastError();              // automatic: fall through to the graph path
  }
  // Sharded steps run fused (exchange inside a rank-sharded lookup) except the LAST step of the
  // call, which exchanges whole histograms: bins / norms / nll_at find them where they always were.
  const bool fused = rebin && chain_fused_ok(c);
  auto plain_step = [&]() -> int {
    const unsigned i = c->steps_done;
    int rc = chain_enqueue_step(c, rebin, rebin ? (int)((i + 1) & 1) : 0, fused && left > 1);
    if (!rc) { c->steps_done += 1; left -= 1; }
    return rc;
  };
  if ((c->mode == 0 || c->mode == 3) && left >= 2 * GRAPH_STEPS && !c->graph_exec) {
    // CUDA graph of GRAPH_STEPS consecutive steps starting at an even step index
    cudaGraph_t graph = nullptr;
    const long before = c->launches;
    SXMC_CUDA(cudaStreamBeginCapture(c->stream, has_comm ? cudaStreamCaptureModeRelaxed
                                                          : cudaStreamCaptureModeThreadLocal));
    int rc = SXMC_OK;
    for (unsigned k = 0; k < GRAPH_STEPS && !rc; ++k) rc = chain_enqueue_step(c, rebin, rebin ? (int)((k + 1) & 1) : 0, fused);
    cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
    c->graph_launches = c->launches - before;
    c->launches = before;
    if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamEndCapture", __FILE__, __LINE__);
    e = cudaGraphInstantiate(&c->graph_exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) { c->graph_exec = nullptr; return cuda_fail(e, "cudaGraphInstantiate", __FILE__, __LINE__); }
  }
  SXMC_CUDA(cudaEventRecord(c->ev0, c->stream));
  if ((c->mode == 0 || c->mode == 3) && c->graph_exec) {
    if ((c->steps_done & 1u) && left > 0) SXMC_TRY(plain_step());
    while (left >= GRAPH_STEPS + (fused ? 1u : 0u)) {
      SXMC_CUDA(cudaGraphLaunch(c->graph_exec, c->stream));
      c->launches += c->graph_launches;
      c->steps_done += GRAPH_STEPS;
      left -= GRAPH_STEPS;
    }
  }
  while (left > 0) SXMC_TRY(plain_step());
  if (rebin && c->steps_done > 0) c->last_buf = (int)(c->steps_done & 1);   // step n-1 used buffer n&1
  SXMC_CUDA(cudaEventRecord(c->ev1, c->stream));
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  float ms = 0;
  SXMC_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
  c->loop_ms += ms;
  c->last_path = 1;
  return chain_p2p_check(c);
}

// Same steps as sxmc_chain_advance, issued as plain stream launches with a CUDA event around
// every kernel, and (flush_bytes > 0) an L2 flush -- a memset of flush_bytes -- before every
// step, outside the timed intervals.  Measurement only.
int sxmc_chain_advance_timed(sxmc_chain* c, unsigned nsteps, size_t flush_bytes, float* step_ms_sum,
                             float* bin_ms_sum, float* nll_ms_sum) {
  if (!c || !c->have_data) { set_error("sxmc_chain_advance_timed: sxmc_chain_begin was not called"); return SXMC_ESTATE; }
  if ((unsigned long long)c->steps_done + nsteps > c->steps_planned) {
    set_error("sxmc_chain_advance_timed: more steps than sxmc_chain_begin announced");
    return SXMC_ESTATE;
  }
  const bool rebin = c->nsystematics > 0 && !c->systematics_fixed;
  SXMC_TRY(chain_fixed_table(c, rebin));
  if (flush_bytes > c->flush_bytes) {
    cudaFree(c->flush_buf);
    c->flush_buf = nullptr;
    c->flush_bytes = 0;
    SXMC_CUDA(cudaMalloc(&c->flush_buf, flush_bytes));
    c->flush_bytes = flush_bytes;
  }
  const bool fused_ok = rebin && chain_fused_ok(c);
  const unsigned CHUNK = 64;
  if (c->tev.empty()) {
    c->tev.resize(4 * CHUNK);
    for (auto& e : c->tev) SXMC_CUDA(cudaEventCreate(&e));
  }
  double t_step = 0, t_bin = 0, t_nll = 0, t_comm = 0;
  unsigned left = nsteps;
  while (left > 0) {
    const unsigned n = std::min(left, CHUNK);
    for (unsigned k = 0; k < n; ++k) {
      const unsigned i = c->steps_done;
      const int b = rebin ? (int)((i + 1) & 1) : 0;
      if (flush_bytes) SXMC_CUDA(cudaMemsetAsync(c->flush_buf, (int)(i & 0xff), flush_bytes, c->stream));
      SXMC_CUDA(cudaEventRecord(c->tev[3 * k], c->stream));
      const bool fused = fused_ok && !(left == n && k + 1 == n);       // not the call's last step
      if (rebin) SXMC_TRY(chain_bin(c, b, c->tev[3 * CHUNK + k], true, fused));
      else SXMC_CUDA(cudaEventRecord(c->tev[3 * CHUNK + k], c->stream));
      SXMC_CUDA(cudaEventRecord(c->tev[3 * k + 1], c->stream));
      SXMC_TRY(chain_nll(c, b, NLL_MODE_STEP, rebin, nullptr, fused));
      SXMC_CUDA(cudaEventRecord(c->tev[3 * k + 2], c->stream));
      c->steps_done += 1;
    }
    SXMC_CUDA(cudaStreamSynchronize(c->stream));
    for (unsigned k = 0; k < n; ++k) {
      float a = 0, b2 = 0;
      SXMC_CUDA(cudaEventElapsedTime(&a, c->tev[3 * k], c->tev[3 * k + 1]));
      SXMC_CUDA(cudaEventElapsedTime(&b2, c->tev[3 * k + 1], c->tev[3 * k + 2]));
      float cm = 0;     // binning kernel end -> all-reduce end (0 without a communicator)
      SXMC_CUDA(cudaEventElapsedTime(&cm, c->tev[3 * CHUNK + k], c->tev[3 * k + 1]));
      t_comm += cm;
      t_bin += a - cm; t_nll += b2; t_step += a + b2;      // bin = the binning launch alone
    }
    left -= n;
  }
  if (rebin && c->steps_done > 0) c->last_buf = (int)(c->steps_done & 1);
  c->loop_ms += (float)t_step;
  if (step_ms_sum) *step_ms_sum = (float)t_step;
  if (bin_ms_sum) *bin_ms_sum = (float)t_bin;
  if (nll_ms_sum) *nll_ms_sum = (float)t_nll;
  c->last_comm_ms = (float)t_comm;
  return chain_p2p_check(c);
}

int sxmc_chain_end(sxmc_chain* c, long* nrows_out) {
  if (!c) SXMC_INVAL("sxmc_chain_end: NULL chain");
  SXMC_TRY(chain_pull_state(c));
  c->nrows = c->st_host->nrows;
  c->accepted = c->st_host->accepted;
  if (nrows_out) *nrows_out = (long)std::min(c->nrows, c->row_cap);
  return SXMC_OK;
}

int sxmc_chain_run(sxmc_chain* c, const float* data, size_t data_nfloats, unsigned nsteps,
                   float burnin_fraction, int debug_mode, unsigned sync_interval, long* nrows_out) {
  (void)sync_interval;
  if (!c) SXMC_INVAL("sxmc_chain_run: NULL chain");
  SXMC_TRY(sxmc_chain_begin(c, data, data_nfloats, nsteps, burnin_fraction, debug_mode));
  SXMC_TRY(sxmc_chain_advance(c, nsteps));
  return sxmc_chain_end(c, nrows_out);
}

int sxmc_chain_read(sxmc_chain* c, float* rows_out, long max_rows) {
  long long n = std::min<long long>(std::min(c->nrows, c->row_cap), max_rows);
  if (n > 0)
    SXMC_CUDA(cudaMemcpy(rows_out, c->rows_dev, sizeof(float) * (size_t)n * (c->nparameters + 1),
                         cudaMemcpyDeviceToHost));
  return SXMC_OK;
}
int sxmc_chain_accepted(sxmc_chain* c, long* accepted_out) { *accepted_out = (long)c->accepted; return SXMC_OK; }
int sxmc_chain_loop_ms(sxmc_chain* c, float* ms_out) { *ms_out = c->loop_ms; return SXMC_OK; }
int sxmc_chain_launches(sxmc_chain* c, long* launches_out) { *launches_out = c->launches; return SXMC_OK; }
int sxmc_chain_last_path(sxmc_chain* c, int* path_out) { *path_out = c->last_path; return SXMC_OK; }
int sxmc_chain_last_comm_ms(sxmc_chain* c, float* ms_out) { *ms_out = c->last_comm_ms; return SXMC_OK; }
int sxmc_chain_last_kernel_ms(sxmc_chain* c, float* bin_ms, float* nll_ms) {
  if (!c) SXMC_INVAL("sxmc_chain_last_kernel_ms: NULL chain");
  if (bin_ms) *bin_ms = c->last_bin_ms;
  if (nll_ms) *nll_ms = c->last_nll_ms;
  return SXMC_OK;
}
int sxmc_chain_bin_kernel_kind(sxmc_chain* c, int* kind_out) {
  if (!c || !kind_out) SXMC_INVAL("sxmc_chain_bin_kernel_kind: NULL argument");
  if (!c->have_data) { set_error("sxmc_chain_bin_kernel_kind: no data set (the launch is planned with the data)"); return SXMC_ESTATE; }
  const BinLaunch& L = c->batch[0].launch;
  *kind_out = !L.fast ? 0 : (L.prog_nops < 0 ? 1 : (L.jit ? 3 : 2));
  return SXMC_OK;
}

int sxmc_chain_bin_plan(sxmc_chain* c, int* nwarps_out, int* rows_per_chunk_out, int* nstages_out, int* smem_hist_out) {
  if (!c) SXMC_INVAL("sxmc_chain_bin_plan: NULL chain");
  if (!c->have_data) { set_error("sxmc_chain_bin_plan: no data set (the launch is planned with the data)"); return SXMC_ESTATE; }
  const BinLaunch& L = c->batch[0].launch;
  if (nwarps_out) *nwarps_out = L.fast ? (L.nwarps > 0 ? L.nwarps : FAST_WARPS) : 0;
  if (rows_per_chunk_out) *rows_per_chunk_out = L.fast ? L.rows_per_stage : 0;
  if (nstages_out) *nstages_out = L.fast ? L.nstages : 0;
  if (smem_hist_out) *smem_hist_out = L.smem_hist ? 1 : 0;
  return SXMC_OK;
}

int sxmc_chain_read_trace(sxmc_chain* c, double* nll_trace_out, double* proposed_out, unsigned nsteps) {
  if (!c->nll_trace_dev || nsteps > c->trace_steps) { set_error("trace was not enabled for that many steps"); return SXMC_ESTATE; }
  if (nll_trace_out) SXMC_CUDA(cudaMemcpy(nll_trace_out, c->nll_trace_dev, sizeof(double) * nsteps, cudaMemcpyDeviceToHost));
  if (proposed_out)
    SXMC_CUDA(cudaMemcpy(proposed_out, c->prop_trace_dev, sizeof(double) * nsteps * c->nparameters, cudaMemcpyDeviceToHost));
  return SXMC_OK;
}

int sxmc_chain_nll_at(sxmc_chain* c, const double* pars, int rebin, double* nll_out, double* event_sum_out) {
  if (!c->have_data) { set_error("sxmc_chain_nll_at: no data set"); return SXMC_ESTATE; }
  if (!c->st_host->nparameters) {   // never run: build a neutral state
    chain_init_state(c, *c->st_host, 0, 0.f, 0);
    SXMC_TRY(chain_push_state(c));
    rebin = 1;
  }
  SXMC_CUDA(cudaMemcpyAsync(c->st_dev->proposed, pars, sizeof(double) * c->nparameters, cudaMemcpyHostToDevice, c->stream));
  const int b = c->last_buf;
  cudaEvent_t e0 = c->ev0, e1 = c->ev1;
  if (rebin) {
    SXMC_CUDA(cudaMemsetAsync(c->part_buf[b], 0, sizeof(unsigned) * c->hist_words, c->stream));
    if (g_timing) SXMC_CUDA(cudaEventRecord(e0, c->stream));
    SXMC_TRY(chain_bin(c, b));
    if (g_timing) {
      SXMC_CUDA(cudaEventRecord(e1, c->stream));
      SXMC_CUDA(cudaEventSynchronize(e1));
      SXMC_CUDA(cudaEventElapsedTime(&c->last_bin_ms, e0, e1));
      g_last_timed[0] = c->last_bin_ms;
    }
  }
  if (g_timing) SXMC_CUDA(cudaEventRecord(e0, c->stream));
  SXMC_TRY(chain_nll(c, b, NLL_MODE_EVAL_ONLY, false, c->lut_dev));
  if (g_timing) {
    SXMC_CUDA(cudaEventRecord(e1, c->stream));
    SXMC_CUDA(cudaEventSynchronize(e1));
    SXMC_CUDA(cudaEventElapsedTime(&c->last_nll_ms, e0, e1));
    g_last_timed[1] = c->last_nll_ms;
  }
  double out[2];
  SXMC_CUDA(cudaMemcpyAsync(out, &c->st_dev->last_nll, sizeof out, cudaMemcpyDeviceToHost, c->stream));
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  if (nll_out) *nll_out = out[0];
  if (event_sum_out) *event_sum_out = out[1];
  return SXMC_OK;
}

int sxmc_chain_get_norms(sxmc_chain* c, unsigned* norms_out) {
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  SXMC_CUDA(cudaMemcpy(norms_out, c->hist_buf[c->last_buf] + (size_t)c->nsignals * c->total_nbins,
                       sizeof(unsigned) * c->nsignals, cudaMemcpyDeviceToHost));
  return SXMC_OK;
}
int sxmc_chain_get_lut(sxmc_chain* c, float* lut_out) {
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  SXMC_CUDA(cudaMemcpy(lut_out, c->lut_dev, sizeof(float) * c->nevents * c->nsignals, cudaMemcpyDeviceToHost));
  return SXMC_OK;
}
int sxmc_chain_get_bins(sxmc_chain* c, unsigned* bins_out) {
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  SXMC_CUDA(cudaMemcpy(bins_out, c->hist_buf[c->last_buf], sizeof(unsigned) * (size_t)c->nsignals * c->total_nbins,
                       cudaMemcpyDeviceToHost));
  return SXMC_OK;
}

// ---------------------------------------------------------------------------------------
// multi-GPU
// ---------------------------------------------------------------------------------------
int sxmc_comm_unique_id(void* id128_out) {
  SXMC_TRY(load_nccl());
  nccl_unique_id id;
  int rc = g_nccl.GetUniqueId(&id);
  if (rc) return nccl_fail(rc, "ncclGetUniqueId");
  memcpy(id128_out, &id, sizeof id);
  return SXMC_OK;
}
int sxmc_comm_create(const void* id128, int nranks, int rank, sxmc_comm** out) {
  SXMC_TRY(ensure_init());
  SXMC_TRY(load_nccl());
  nccl_unique_id id;
  memcpy(&id, id128, sizeof id);
  sxmc_comm* c = new sxmc_comm;
  c->nranks = nranks; c->rank = rank;
  int rc = g_nccl.CommInitRank(&c->comm, nranks, id, rank);
  if (rc) { delete c; return nccl_fail(rc, "ncclCommInitRank"); }
  *out = c;
  return SXMC_OK;
}
int sxmc_comm_destroy(sxmc_comm* comm) {
  if (!comm) return SXMC_OK;
  if (comm->comm) g_nccl.CommDestroy(comm->comm);
  delete comm;
  return SXMC_OK;
}
int sxmc_comm_allreduce_u32(sxmc_comm* comm, unsigned* dev, size_t count) {
  int rc = g_nccl.AllReduce(dev, dev, count, NCCL_UINT32, NCCL_SUM, comm->comm, 0);
  if (rc) return nccl_fail(rc, "ncclAllReduce");
  SXMC_CUDA(cudaStreamSynchronize(0));
  return SXMC_OK;
}

}  // extern "C"


// =========================================================================================
// Sample sharding over NVLink peer memory: the chain's partial histograms live in a window the
// other ranks open with CUDA IPC; the per-step exchange is hist_allreduce_p2p_kernel.
// =========================================================================================
extern "C" {

int sxmc_chain_p2p_export(sxmc_chain* c, void* handle64_out) {
  if (!c || !handle64_out) SXMC_INVAL("sxmc_chain_p2p_export: NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  if (!c->p2p_window) {
    c->p2p_pitch = (c->hist_words + 31) & ~(size_t)31;           // 128-byte aligned parities
    // ... + [P2P_PARTIALS_MAX] x 16 bytes: the tagged partial sums of the fused, rank-sharded lookup
    const size_t words = 4 * c->p2p_pitch + 2 * P2P_MAX_RANKS + 4 * (size_t)P2P_PARTIALS_MAX;
    SXMC_CUDA(cudaMalloc(&c->p2p_window, sizeof(unsigned) * words));
    SXMC_CUDA(cudaMemset(c->p2p_window, 0, sizeof(unsigned) * words));
    SXMC_CUDA(cudaMalloc(&c->p2p_state, sizeof(unsigned) * 32));      // + 8 x u64 phase clocks
    SXMC_CUDA(cudaMemset(c->p2p_state, 0, sizeof(unsigned) * 32));
    SXMC_CUDA(cudaMallocHost(&c->p2p_state_host, sizeof(unsigned) * 32));
  }
  cudaIpcMemHandle_t h;
  SXMC_CUDA(cudaIpcGetMemHandle(&h, c->p2p_window));
  memcpy(handle64_out, &h, 64);
  return SXMC_OK;
}

int sxmc_chain_p2p_connect(sxmc_chain* c, const void* handles, int nranks, int rank) {
  if (!c || !handles) SXMC_INVAL("sxmc_chain_p2p_connect: NULL argument");
  if (nranks < 1 || nranks > P2P_MAX_RANKS || rank < 0 || rank >= nranks)
    SXMC_INVAL("sxmc_chain_p2p_connect: 1..8 ranks of one box");
  if (!c->p2p_window) { set_error("sxmc_chain_p2p_connect: call sxmc_chain_p2p_export first"); return SXMC_ESTATE; }
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  chain_drop_graph(c);
  for (int p = 0; p < nranks; ++p) {
    if (p == rank) { c->p2p_peer[p] = c->p2p_window; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)handles + 64 * (size_t)p, 64);
    SXMC_CUDA(cudaIpcOpenMemHandle(&c->p2p_peer[p], h, cudaIpcMemLazyEnablePeerAccess));
  }
  c->p2p_nranks = nranks;
  c->p2p_rank = rank;
  if (const char* ts = std::getenv("SXMC_P2P_TIMEOUT_S")) {
    const double sec = std::atof(ts);
    if (sec > 0) c->p2p_timeout_ns = (unsigned long long)(sec * 1e9);
  }
  c->part_buf[0] = c->p2p_window;
  c->part_buf[1] = c->p2p_window + c->p2p_pitch;
  c->hist_buf[0] = c->p2p_window + 2 * c->p2p_pitch;      // peers store their slices of the sum here
  c->hist_buf[1] = c->p2p_window + 3 * c->p2p_pitch;
  c->table_sig.clear();
  if (c->have_data) SXMC_TRY(chain_build_tables(c));
  return SXMC_OK;
}

int sxmc_chain_p2p_status(sxmc_chain* c, unsigned* exchanges_out, int* aborted_out) {
  if (!c || !c->p2p_state) { if (exchanges_out) *exchanges_out = 0; if (aborted_out) *aborted_out = 0; return SXMC_OK; }
  SXMC_CUDA(cudaMemcpy(c->p2p_state_host, c->p2p_state, sizeof(unsigned) * 4, cudaMemcpyDeviceToHost));
  if (exchanges_out) *exchanges_out = c->p2p_state_host[0];
  if (aborted_out) *aborted_out = (int)c->p2p_state_host[2];
  return SXMC_OK;
}

int sxmc_chain_p2p_clocks(sxmc_chain* c, double* ns_out7) {
  if (!c || !ns_out7) SXMC_INVAL("sxmc_chain_p2p_clocks: NULL argument");
  for (int i = 0; i < 7; ++i) ns_out7[i] = 0;
  if (!c->p2p_state) return SXMC_OK;
  SXMC_CUDA(cudaStreamSynchronize(c->stream));
  SXMC_CUDA(cudaMemcpy(c->p2p_state_host, c->p2p_state, sizeof(unsigned) * 32, cudaMemcpyDeviceToHost));
  const unsigned long long* clk = reinterpret_cast<const unsigned long long*>(c->p2p_state_host + 4);
  for (int i = 0; i < 7; ++i) ns_out7[i] = (double)clk[1 + i];
  return SXMC_OK;
}

}  // extern "C"
