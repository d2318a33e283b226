// Pointwise (1x1) convolution on channels-last activations as a tcgen05 GEMM whose epilogue
// also produces the per-channel sum and sum of squares BatchNorm needs.
//
//   Y[M, N] = X[M, K] * W[N, K]^T        M = batch*H*W pixels, K = C_in, N = C_out   (bf16)
//   sum[n]  = sum_m acc[m, n]            fp32, from the fp32 accumulators
//   sqs[n]  = sum_m acc[m, n]^2
//
// Why: after the BatchNorm / ReLU / pooling kernels of batchnorm.cu the ResNet-50 step still
// reads every convolution output once just to take its statistics (`bn_stats`, 1.45 ms of a
// 25.2 ms step, profiles/README.md §7); five sixths of those bytes come out of 1x1
// convolutions, which are plain GEMMs on NHWC memory. Taking the statistics while the tile is
// still in tensor memory removes that pass.
//
// How: the mainloop is CUTLASS's sm_100 warp-specialised collective (TMA loads into a
// shared-memory ring, one elected thread issuing tcgen05.mma with the accumulator in TMEM); the
// epilogue is an epilogue-visitor tree written here: two row-vector reductions over the
// accumulator fragment (x and x*x) that finish with fp32 atomics into sum / sqs, and the bf16
// store of Y through TMA. The statistics are therefore summed in a run-dependent order (last
// bits may differ between runs); BatchNorm statistics are per rank and never compared across
// ranks, so replica consistency is not affected.
//
// STATUS: compiled for sm_100a and inspected (SASS shows UTCMMA / UTMALDG / UTMASTG); it has not
// run on a GPU yet (round-1 GPU time was used up) and nothing calls it by default.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <string>

#include "fmpi.h"

#include "cutlass/cutlass.h"
#include "cutlass/bfloat16.h"
#include "cutlass/functional.h"
#include "cutlass/epilogue/collective/collective_builder.hpp"
#include "cutlass/epilogue/fusion/sm90_visitor_compute_tma_warpspecialized.hpp"
#include "cutlass/epilogue/fusion/sm90_visitor_load_tma_warpspecialized.hpp"
#include "cutlass/epilogue/fusion/sm90_visitor_store_tma_warpspecialized.hpp"
#include "cutlass/epilogue/fusion/sm90_visitor_tma_warpspecialized.hpp"
#include "cutlass/gemm/collective/collective_builder.hpp"
#include "cutlass/gemm/device/gemm_universal_adapter.h"
#include "cutlass/gemm/kernel/gemm_universal.hpp"
#include "cutlass/util/packed_stride.hpp"

namespace {

thread_local std::string g_gemm_err;

int gemm_fail(const char* what, const char* detail) {
  g_gemm_err = std::string(what) + ": " + detail;
  return 1;
}

using namespace cute;
namespace fus = cutlass::epilogue::fusion;

// f(a, b) = a: lets the visitor tree evaluate a second branch (the x*x reduction) for its side
// effect while the stored value stays the accumulator.
template <class T>
struct keep_first {
  CUTLASS_HOST_DEVICE T operator()(T const& a, T const&) const { return a; }
};

constexpr auto kRound = cutlass::FloatRoundStyle::round_to_nearest;

// k2Sm = false: one SM per 128 x kTileN tile (CTA tile == MMA tile).
// k2Sm = true:  the CTA pair of a cluster shares one 256 x kTileN MMA tile (tcgen05
//               cta_group::2): each CTA still owns a 128 x kTileN accumulator and epilogue, the W
//               tile is loaded once for both.
template <int kTileN, bool k2Sm>
struct PointwiseConvStats {
  using Element = cutlass::bfloat16_t;
  using CtaTile = Shape<_128, Int<kTileN>, _64>;
  using TileShape = cute::conditional_t<k2Sm, Shape<_256, Int<kTileN>, _64>, CtaTile>;   // MMA tile
  using ClusterShape = cute::conditional_t<k2Sm, Shape<_2, _1, _1>, Shape<_1, _1, _1>>;
  using EpilogueSchedule = cute::conditional_t<k2Sm, cutlass::epilogue::TmaWarpSpecialized2Sm,
                                               cutlass::epilogue::TmaWarpSpecialized1Sm>;
  using MainloopSchedule = cute::conditional_t<k2Sm, cutlass::gemm::KernelTmaWarpSpecialized2SmSm100,
                                               cutlass::gemm::KernelTmaWarpSpecialized1SmSm100>;
  static constexpr int kAlign = 128 / cutlass::sizeof_bits<Element>::value;

  using RowSum = fus::Sm90RowReduction<cutlass::plus, cutlass::plus, cutlass::atomic_add, 0, CtaTile,
                                       float, float, kRound>;
  using Square = fus::Sm90EVT<fus::Sm90Compute<cutlass::multiplies, float, float, kRound>,
                              fus::Sm90AccFetch, fus::Sm90AccFetch>;
  using SumBranch = fus::Sm90EVT<RowSum, fus::Sm90AccFetch>;      // passes the accumulator on
  using SqsBranch = fus::Sm90EVT<RowSum, Square>;                 // passes its square on
  using Tree = fus::Sm90EVT<fus::Sm90Compute<keep_first, Element, float, kRound>, SumBranch, SqsBranch>;

  using Epilogue = typename cutlass::epilogue::collective::CollectiveBuilder<
      cutlass::arch::Sm100, cutlass::arch::OpClassTensorOp, TileShape, ClusterShape,
      cutlass::epilogue::collective::EpilogueTileAuto, float, float,
      void, cutlass::layout::RowMajor, kAlign,                    // no source tensor
      Element, cutlass::layout::RowMajor, kAlign,
      EpilogueSchedule, Tree>::CollectiveOp;

  using Mainloop = typename cutlass::gemm::collective::CollectiveBuilder<
      cutlass::arch::Sm100, cutlass::arch::OpClassTensorOp,
      Element, cutlass::layout::RowMajor, kAlign,                 // X: [M, K], K contiguous
      Element, cutlass::layout::ColumnMajor, kAlign,              // W: [N, K], K contiguous
      float, TileShape, ClusterShape,
      cutlass::gemm::collective::StageCountAutoCarveout<static_cast<int>(sizeof(typename Epilogue::SharedStorage))>,
      MainloopSchedule>::CollectiveOp;

  using Kernel = cutlass::gemm::kernel::GemmUniversal<Shape<int, int, int, int>, Mainloop, Epilogue>;
  using Gemm = cutlass::gemm::device::GemmUniversalAdapter<Kernel>;

  static typename Gemm::Arguments arguments(const void* x, const void* w, void* y, float* sum, float* sqs,
                                            int M, int N, int K) {
    using StrideA = typename Kernel::StrideA;
    using StrideB = typename Kernel::StrideB;
    using StrideD = typename Kernel::StrideD;
    auto sa = cutlass::make_cute_packed_stride(StrideA{}, make_shape(M, K, 1));
    auto sb = cutlass::make_cute_packed_stride(StrideB{}, make_shape(N, K, 1));
    auto sd = cutlass::make_cute_packed_stride(StrideD{}, make_shape(M, N, 1));
    typename Tree::Arguments tree{
        {{}, {sum, 0.0f, {}}},              // SumBranch: AccFetch, RowSum
        {{{}, {}, {}}, {sqs, 0.0f, {}}},    // SqsBranch: Square(AccFetch, AccFetch, multiplies), RowSum
        {}};                                // keep_first
    return typename Gemm::Arguments{
        cutlass::gemm::GemmUniversalMode::kGemm,
        {M, N, K, 1},
        {static_cast<const Element*>(x), sa, static_cast<const Element*>(w), sb},
        {tree, nullptr, sd, static_cast<Element*>(y), sd}};
  }

  static size_t workspace_bytes(int M, int N, int K) {
    Gemm gemm;
    return gemm.get_workspace_size(arguments(nullptr, nullptr, nullptr, nullptr, nullptr, M, N, K));
  }

  static int run(const void* x, const void* w, void* y, float* sum, float* sqs, int M, int N, int K,
                 void* workspace, size_t workspace_size, cudaStream_t st) {
    Gemm gemm;
    auto args = arguments(x, w, y, sum, sqs, M, N, K);
    if (gemm.get_workspace_size(args) > workspace_size) return gemm_fail("gemm_bnstats", "workspace too small");
    cutlass::Status s = gemm.can_implement(args);
    if (s != cutlass::Status::kSuccess) return gemm_fail("gemm_bnstats can_implement", cutlassGetStatusString(s));
    s = gemm.initialize(args, workspace, st);     // also zeroes sum / sqs on the stream
    if (s != cutlass::Status::kSuccess) return gemm_fail("gemm_bnstats initialize", cutlassGetStatusString(s));
    s = gemm.run(st);
    if (s != cutlass::Status::kSuccess) return gemm_fail("gemm_bnstats run", cutlassGetStatusString(s));
    return 0;
  }
};

}  // namespace

extern "C" {

const char* fmpi_gemm_last_error(void) { return g_gemm_err.c_str(); }

// CTA tiles are 128 x 64 x 64: every thread of the four epilogue warps then keeps 64 + 64
// partial column sums in registers (a 128-wide tile needs 256 and spills; see -Xptxas -v).
// The layers this serves are bandwidth-bound, so the narrower tile costs L2 re-reads of X only.
// Two variants to time against each other: 0 = one SM per tile, 1 = CTA pairs (cta_group::2).
using ConvStats1Sm = PointwiseConvStats<64, false>;
using ConvStats2Sm = PointwiseConvStats<64, true>;
static int g_variant = 0;

int fmpi_conv1x1_bnstats_variant(int variant) {
  if (variant < 0 || variant > 1) return gemm_fail("fmpi_conv1x1_bnstats_variant", "0 (1 SM) or 1 (2 SM)");
  g_variant = variant;
  return 0;
}

// Bytes of device workspace fmpi_conv1x1_bnstats needs for this problem (may be 0).
size_t fmpi_conv1x1_bnstats_workspace(int M, int N, int K) {
  return g_variant ? ConvStats2Sm::workspace_bytes(M, N, K) : ConvStats1Sm::workspace_bytes(M, N, K);
}

// y[M, N] = x[M, K] * w[N, K]^T (all bf16, row-major, 16-byte aligned rows: K % 8 == 0 and
// N % 8 == 0); sum[N] and sqs[N] (fp32) receive the per-column sum and sum of squares of the
// fp32 accumulators. Stream-ordered; sum / sqs are zeroed by the call itself.
int fmpi_conv1x1_bnstats(const void* x, const void* w, void* y, float* sum, float* sqs, int M, int N, int K,
                         void* workspace, size_t workspace_bytes, void* stream) {
  if (!x || !w || !y || !sum || !sqs) return gemm_fail("fmpi_conv1x1_bnstats", "null argument");
  if (M <= 0 || N <= 0 || K <= 0 || (N % 8) || (K % 8))
    return gemm_fail("fmpi_conv1x1_bnstats", "need M, N, K > 0 with N and K multiples of 8");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return g_variant ? ConvStats2Sm::run(x, w, y, sum, sqs, M, N, K, workspace, workspace_bytes, st)
                   : ConvStats1Sm::run(x, w, y, sum, sqs, M, N, K, workspace, workspace_bytes, st);
}

}  // extern "C"
