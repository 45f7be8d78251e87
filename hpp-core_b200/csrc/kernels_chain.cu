// PathVector recursion of ContinuousValidation::validate (reference src/continuous-validation.cc:153-192) for chains of
// StraightPath leaves.  The reference validates the sub-paths one after the other and stops at the first invalid one;
// the sub-path validations are independent, so the engine validates ALL leaves of ALL chains as one edge batch
// (hppb200_validate_edges machinery) and this kernel replays the recursion's bookkeeping per chain: which leaf fails
// first in walking order, the parameter offset added to its report (`param`, accumulated leaf by leaf exactly like
// the reference: += length() going forward, -= length() going backward from PathVector::length()), where the valid
// part ends.  Results of leaves behind the first failure are simply not looked at.
#include <math_constants.h>

#include "engine.hpp"

namespace hppb200 {

__global__ void chainReduceKernel(const int32_t* __restrict__ chainBegin, size_t nChains, int reverse, const double* __restrict__ L,
                                  const uint8_t* __restrict__ eValid, const double* __restrict__ eTfail,
                                  const double* __restrict__ eTlast, const int32_t* __restrict__ ePair,
                                  const uint8_t* __restrict__ eReason, uint8_t* valid, double* tFail, double* tLast, int32_t* pair,
                                  uint8_t* reason, int32_t* firstBad) {
  const size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= nChains) return;
  const int b = chainBegin[c], e = chainBegin[c + 1];
  int bad = -1;
  double param = 0, tf = CUDART_NAN, tl = 0;
  if (!reverse) {
    for (int i = b; i < e; ++i) {
      if (eValid[i]) param += L[i];  // :183-185
      else { bad = i; tf = eTfail[i] + param; tl = param + eTlast[i]; break; }  // :186-189
    }
    if (bad < 0) tl = param;
  } else {
    for (int i = b; i < e; ++i) param += L[i];  // PathVector::length(): timeRange accumulated by appendPath
    for (int i = e - 1; i >= b; --i) {
      if (eValid[i]) param -= L[i];  // :163-165
      else { bad = i; tf = eTfail[i] + (param - L[i]); tl = (param - L[i]) + eTlast[i]; break; }  // :166-174
    }
    if (bad < 0) tl = 0.0;
  }
  valid[c] = bad < 0 ? 1 : 0;
  if (tFail) tFail[c] = tf;
  if (tLast) tLast[c] = tl;
  if (pair) pair[c] = bad < 0 ? -1 : ePair[bad];
  if (reason) reason[c] = bad < 0 ? (uint8_t)HPPB200_REASON_VALID : eReason[bad];
  if (firstBad) firstBad[c] = bad;
}

void launchChainReduce(Workspace* ws, const int32_t* d_chainBegin, size_t nChains, int reverse, const double* d_L, const uint8_t* eValid,
                       const double* eTfail, const double* eTlast, const int32_t* ePair, const uint8_t* eReason, uint8_t* valid,
                       double* tFail, double* tLast, int32_t* pair, uint8_t* reason, int32_t* firstBad) {
  if (!nChains) return;
  chainReduceKernel<<<(unsigned)((nChains + 127) / 128), 128, 0, ws->stream>>>(d_chainBegin, nChains, reverse, d_L, eValid, eTfail, eTlast,
                                                                              ePair, eReason, valid, tFail, tLast, pair, reason, firstBad);
  HPPB200_CUDA_CHECK(cudaGetLastError());
  ws->stats.kernel_launches += 1;
}

}  // namespace hppb200
