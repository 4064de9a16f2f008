// tools/intpipe_bench.cu -- issue-rate microbenchmark of the SM integer pipes on sm_100a.
//
// K2's binding resource is not HBM but the ALU pipe that executes VABSDIFF4 (the 4-byte sum of absolute differences the
// reference's FPGA kernel does with a 32-lane abs-diff + adder tree, hls_opencl/kernel/myMatch.cpp:159-178).  This
// program measures, per SM and clock, how many warp instructions of each kind the machine sustains, alone and in the
// mixes K2's inner loops issue, so that "ALU pipe N %" becomes a fraction of a MEASURED integer roofline:
//
//     peak SAD32/s = (VABSDIFF4 warp-instr / clk / SM) * 32 lanes / 8 instr per 32-byte SAD * SMs * SM clock
//
// Build + run (one GPU, ~2 s):
//     nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/intpipe_bench tools/intpipe_bench.cu
//     tools/intpipe_bench > gpurun_out/intpipe.json
// Every loop body is inline PTX on 8 independent accumulator chains per thread (latency 4 x rate 2 needs >= 2 warps x 4
// chains per scheduler; 8 warps x 8 chains leave no dependency stall), 32 warps per SM resident.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int ITERS = 8192;   // loop trips (~0.6 ms per measurement: SM clocks settle)
constexpr int CHAINS = 8;     // independent accumulators per thread
constexpr int THREADS = 256;
constexpr int CTAS_PER_SM = 4;

enum Op { VABS_ACC, VABS_NOACC, IADD3, LOP3, VIMNMX, SHF, IMAD, DP4A, PRMT, POPC, CLZ, BALLOT, REDUX_MIN, SHFL, LDS128_BCAST, LDS32_BCAST,
          LDS32_GATHER, MIX_VABS_IMAD, MIX_VABS_DP4A, MIX_SAD32_KEY, MIX_FILTER2, IMADHI, MIX_F3_SHF, MIX_F3_MADHI, MIX_F3_ADDC, MIX_F3_MADHI1, N_OPS };
static const char* kNames[N_OPS] = {
    "vabsdiff4_acc", "vabsdiff4", "iadd3", "lop3", "vimnmx_u32", "shf_l_wrap", "imad", "dp4a_u32", "prmt", "popc", "clz", "ballot",
    "redux_min_u32", "shfl_idx", "lds128_broadcast", "lds32_broadcast", "lds32_gather_conflict_free", "mix_1vabs_1imad", "mix_1vabs_1dp4a",
    "mix_sad32_key(8vabs+imad+min)", "mix_filter(2vabs+and+shf per cand)",
    "imad_hi_u32", "mix_filter3(2vabs+shf)", "mix_filter4(2vabs+mad.hi+mad.lo)", "mix_filter4(2vabs+add.cc+addc)", "mix_filter3(2vabs+mad.hi count)"};
// warp instructions per chain step (for the rate arithmetic)
static const int kInstPerStep[N_OPS] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 10, 4, 1, 3, 4, 4, 3};

template <int OP>
__global__ void __launch_bounds__(THREADS) bench(uint32_t* out, long long* cycles, uint32_t seed) {
  __shared__ uint4 sm[64];
  __shared__ long long t_first, t_last;
  if (threadIdx.x == 0) { t_first = 0x7FFFFFFFFFFFFFFFLL; t_last = 0; }
  uint32_t a[CHAINS], b = seed ^ threadIdx.x, c = seed * 2654435761u + 12345u;
#pragma unroll
  for (int i = 0; i < CHAINS; i++) a[i] = seed + i * 0x01020304u + threadIdx.x;
  if (threadIdx.x < 64) sm[threadIdx.x] = make_uint4(seed, seed + 1, seed + 2, seed + threadIdx.x);
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < CHAINS; i++) {
      if (OP == VABS_ACC) asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
      if (OP == VABS_NOACC) asm volatile("vabsdiff4.u32.u32.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
      if (OP == IADD3) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
      if (OP == LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
      if (OP == VIMNMX) asm volatile("min.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b + i));
      if (OP == SHF) asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(a[i]) : "r"(b));
      if (OP == IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
      if (OP == DP4A) asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
      if (OP == PRMT) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c & 0x7777u));
      if (OP == POPC) asm volatile("popc.b32 %0, %0;" : "+r"(a[i]));
      if (OP == CLZ) asm volatile("clz.b32 %0, %0;" : "+r"(a[i]));
      if (OP == BALLOT) asm volatile("{ .reg .pred p; setp.ne.u32 p, %0, 0; vote.sync.ballot.b32 %0, p, 0xffffffff; }" : "+r"(a[i]));
      if (OP == REDUX_MIN) asm volatile("redux.sync.min.u32 %0, %0, 0xffffffff;" : "+r"(a[i]));
      if (OP == SHFL) asm volatile("shfl.sync.idx.b32 %0, %0, %1, 0x1f, 0xffffffff;" : "+r"(a[i]) : "r"(b & 31u));
      if (OP == LDS128_BCAST) {
        uint4 v = sm[(it + i) & 63];  // warp-uniform address: one broadcast wavefront
        a[i] ^= v.x ^ v.w;
      }
      if (OP == LDS32_BCAST) a[i] ^= reinterpret_cast<const uint32_t*>(sm)[(it + i) & 255];
      if (OP == LDS32_GATHER) a[i] ^= reinterpret_cast<const uint32_t*>(sm)[(threadIdx.x + it + i) & 255];  // 32 distinct banks
      if (OP == MIX_VABS_IMAD) {
        asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[(i + 4) & 7]) : "r"(b), "r"(c));
      }
      if (OP == MIX_VABS_DP4A) {
        asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
        asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[(i + 4) & 7]) : "r"(b), "r"(c));
      }
      if (OP == MIX_SAD32_KEY) {  // one exact SAD of K2: 8 VABSDIFF4.ACC, key = sad * 2^19 + rank (IMAD), running min (VIMNMX)
        uint32_t s = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(s) : "r"(a[i]), "r"(c + k));
        asm volatile("mad.lo.u32 %0, %0, 0x80000, %1;" : "+r"(s) : "r"(c));
        asm volatile("min.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(s));
      }
      if (OP == IMADHI) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
      if (OP == MIX_F3_SHF) {     // per-candidate sign collection, all on the ALU pipe
        uint32_t s = 0x7FFFFF00u;
        asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(s) : "r"(b), "r"(a[i]));
        asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(s) : "r"(c), "r"(a[i]));
        asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(a[i]) : "r"(s));
      }
      if (OP == MIX_F3_MADHI) {   // sign bit through the FMA pipe: t = (s * 2) >> 32, m = m * 2 + t
        uint32_t s = 0x7FFFFF00u, t;
        asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(s) : "r"(b), "r"(a[i]));
        asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(s) : "r"(c), "r"(a[i]));
        asm volatile("mad.hi.u32 %0, %1, 2, 0;" : "=r"(t) : "r"(s));
        asm volatile("mad.lo.u32 %0, %0, 2, %1;" : "+r"(a[i]) : "r"(t));
      }
      if (OP == MIX_F3_ADDC) {    // sign bit through the carry: s + s sets CF, m = m + m + CF
        uint32_t s = 0x7FFFFF00u, t;
        asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(s) : "r"(b), "r"(a[i]));
        asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(s) : "r"(c), "r"(a[i]));
        asm volatile("add.cc.u32 %0, %1, %1;\n\taddc.u32 %2, %2, %2;" : "=r"(t), "+r"(s), "+r"(a[i]));
      }
      if (OP == MIX_F3_MADHI1) {  // hit COUNT only: m += (s * 2) >> 32
        uint32_t s = 0x7FFFFF00u;
        asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(s) : "r"(b), "r"(a[i]));
        asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(s) : "r"(c), "r"(a[i]));
        asm volatile("mad.hi.u32 %0, %1, 2, %0;" : "+r"(a[i]) : "r"(s));
      }
      if (OP == MIX_FILTER2) {    // one candidate of K2's two-word filter: bias + Q0 + Q1, sign collected by AND and a funnel shift
        uint32_t s = 0x7FFFFF00u;
        asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(s) : "r"(b), "r"(a[i]));
        asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(s) : "r"(c), "r"(a[i]));
        asm volatile("and.b32 %0, %0, %1;" : "+r"(s) : "r"(b));
        asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(a[i]) : "r"(s));
      }
    }
  }
  const long long t1 = clock64();
  uint32_t r = 0;
#pragma unroll
  for (int i = 0; i < CHAINS; i++) r ^= a[i];
  out[blockIdx.x * THREADS + threadIdx.x] = r;
  // span of the CTA on its SM's clock: first warp in to last warp out (warps of a CTA do not finish together)
  if ((threadIdx.x & 31) == 0) { atomicMin(&t_first, t0); atomicMax(&t_last, t1); }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t smid;
    asm("mov.u32 %0, %%smid;" : "=r"(smid));
    cycles[blockIdx.x * 3 + 0] = t_first; cycles[blockIdx.x * 3 + 1] = t_last; cycles[blockIdx.x * 3 + 2] = smid;
  }
}

template <int OP>
static void run(int sms, uint32_t* d_out, long long* d_cyc, long long* h_cyc, double clk_ghz, bool first) {
  const int grid = sms * CTAS_PER_SM;
  bench<OP><<<grid, THREADS>>>(d_out, d_cyc, 1u);  // warm-up
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  CK(cudaEventRecord(e0));
  bench<OP><<<grid, THREADS>>>(d_out, d_cyc, 7u);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  CK(cudaMemcpy(h_cyc, d_cyc, sizeof(long long) * grid * 3, cudaMemcpyDeviceToHost));
  // per SM: clocks from the first warp in to the last warp out over the CTAs that ran there, and the work they did
  static long long sm_first[1024], sm_last[1024]; static int ctas[1024];
  for (int i = 0; i < 1024; i++) { sm_first[i] = 0x7FFFFFFFFFFFFFFFLL; sm_last[i] = 0; ctas[i] = 0; }
  for (int i = 0; i < grid; i++) {
    const int smid = (int)h_cyc[i * 3 + 2] & 1023;
    if (h_cyc[i * 3] < sm_first[smid]) sm_first[smid] = h_cyc[i * 3];
    if (h_cyc[i * 3 + 1] > sm_last[smid]) sm_last[smid] = h_cyc[i * 3 + 1];
    ctas[smid]++;
  }
  double cyc = 0, rate = 0; int used = 0;
  for (int i = 0; i < 1024; i++) if (ctas[i]) {
    const double span = (double)(sm_last[i] - sm_first[i]);
    rate += (double)ITERS * CHAINS * kInstPerStep[OP] * (THREADS / 32) * ctas[i] / span;
    cyc += span; used++;
  }
  cyc /= used;
  const double per_clk_sm = rate / used;
  const double eff_ghz = cyc / (ms * 1e6);  // SM clock implied by CUDA events vs clock64
  printf("%s\n  {\"op\": \"%s\", \"warp_inst_per_clk_per_sm\": %.3f, \"ms\": %.4f, \"sm_clock_ghz_implied\": %.3f}", first ? "" : ",",
         kNames[OP], per_clk_sm, ms, eff_ghz);
  (void)clk_ghz;
}

int main() {
  int dev = 0, sms = 0, khz = 0;
  CK(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, dev));
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
  uint32_t* d_out; long long* d_cyc;
  const int grid = sms * CTAS_PER_SM;
  CK(cudaMalloc(&d_out, sizeof(uint32_t) * grid * THREADS));
  CK(cudaMalloc(&d_cyc, sizeof(long long) * grid * 3));
  long long* h_cyc = (long long*)malloc(sizeof(long long) * grid * 3);
  const double ghz = khz * 1e-6;
  printf("{\"device\": \"%s\", \"sms\": %d, \"sm_clock_max_ghz\": %.3f, \"threads_per_sm\": %d, \"results\": [", prop.name, sms, ghz,
         THREADS * CTAS_PER_SM);
  run<VABS_ACC>(sms, d_out, d_cyc, h_cyc, ghz, true);
  run<VABS_NOACC>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<IADD3>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<LOP3>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<VIMNMX>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<SHF>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<IMAD>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<DP4A>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<PRMT>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<POPC>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<CLZ>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<BALLOT>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<REDUX_MIN>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<SHFL>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<LDS128_BCAST>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<LDS32_BCAST>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<LDS32_GATHER>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<MIX_VABS_IMAD>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<MIX_VABS_DP4A>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<MIX_SAD32_KEY>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<MIX_FILTER2>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<IMADHI>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<MIX_F3_SHF>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<MIX_F3_MADHI>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<MIX_F3_ADDC>(sms, d_out, d_cyc, h_cyc, ghz, false);
  run<MIX_F3_MADHI1>(sms, d_out, d_cyc, h_cyc, ghz, false);
  printf("\n]}\n");
  return 0;
}
