// ffi_evoxmo.cc - XLA FFI handlers over the C ABI of libevoxmo.so (include/evoxmo.h), for the in-jit
// drop-in of EvoX's operators (`jax.ffi.ffi_call` targets).
//
// Built ONLY where JAX is installed: evox_b200/build.py compiles this file into libevoxmo_ffi.so when
// `jax.ffi.include_dir()` exists (g++ -I<that dir> -I include -shared -fPIC ... -L evox_b200 -levoxmo).
// JAX is absent from the build image and from the GPU box this repo is tested on, so this file is
// compiled and exercised by neither; the ctypes + device-pointer binding (evox_b200/) is the tested one.
// Every handler enqueues on XLA's stream and returns; none synchronises, so they are usable inside
// jit / lax.cond / fori_loop.  Scratch memory is a second result buffer that XLA allocates (sizes come
// from the *_workspace_bytes functions at trace time, shapes being static under jit).
#include <cuda_runtime_api.h>

#include "evoxmo.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

static ffi::Error status(int rc) {
    return rc == EVX_OK ? ffi::Error::Success() : ffi::Error(ffi::ErrorCode::kInternal, evx_last_error());
}

// non_dominated_sort(x) -> rank                                   operators/selection/non_dominate.py:32-115
static ffi::Error NdsImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> x, ffi::ResultBuffer<ffi::S32> rank,
                          ffi::ResultBuffer<ffi::U8> ws) {
    auto d = x.dimensions();
    return status(evx_nds_f32(x.typed_data(), d[0], (int32_t)d[1], rank->typed_data(), ws->typed_data(),
                              ws->element_count(), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(EvxNds, NdsImpl,
                              ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>().Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::S32>>().Ret<ffi::Buffer<ffi::U8>>());

// crowding_distance(costs, mask) -> dist                           non_dominate.py:118-158
static ffi::Error CrowdingImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> costs, ffi::Buffer<ffi::PRED> mask,
                               ffi::ResultBuffer<ffi::F32> dist, ffi::ResultBuffer<ffi::U8> ws) {
    auto d = costs.dimensions();
    return status(evx_crowding_distance_f32(costs.typed_data(), reinterpret_cast<const uint8_t*>(mask.untyped_data()), d[0],
                                            (int32_t)d[1], dist->typed_data(), ws->typed_data(), ws->element_count(), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(EvxCrowding, CrowdingImpl,
                              ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>().Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::PRED>>().Ret<ffi::Buffer<ffi::F32>>().Ret<ffi::Buffer<ffi::U8>>());

// crowding_distance_sort(x, mask) -> order                         non_dominate.py:161-180
static ffi::Error CrowdingSortImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> costs, ffi::Buffer<ffi::PRED> mask,
                                   ffi::ResultBuffer<ffi::S32> order, ffi::ResultBuffer<ffi::U8> ws) {
    auto d = costs.dimensions();
    return status(evx_crowding_distance_sort_f32(costs.typed_data(), reinterpret_cast<const uint8_t*>(mask.untyped_data()),
                                                 d[0], (int32_t)d[1], order->typed_data(), ws->typed_data(),
                                                 ws->element_count(), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(EvxCrowdingSort, CrowdingSortImpl,
                              ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>().Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::PRED>>().Ret<ffi::Buffer<ffi::S32>>().Ret<ffi::Buffer<ffi::U8>>());

// non_dominate(population, fitness, topk, deduplicate=False) -> (population[topk], fitness[topk])   :210-222
static ffi::Error SelectImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> pop, ffi::Buffer<ffi::F32> fit,
                             ffi::ResultBuffer<ffi::F32> out_pop, ffi::ResultBuffer<ffi::F32> out_fit,
                             ffi::ResultBuffer<ffi::U8> ws) {
    auto dp = pop.dimensions();
    auto df = fit.dimensions();
    const int64_t topk = out_fit->dimensions()[0];
    return status(evx_non_dominate_f32(pop.typed_data(), fit.typed_data(), nullptr, df[0], (int32_t)dp[1], (int32_t)df[1], topk,
                                       out_pop->typed_data(), out_fit->typed_data(), nullptr, ws->typed_data(),
                                       ws->element_count(), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(EvxNonDominate, SelectImpl,
                              ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>().Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>().Ret<ffi::Buffer<ffi::F32>>().Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::U8>>());

// dominate_relation(x, y) -> bool [nx, ny]                         utils/common.py:94-97
static ffi::Error RelationImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> x, ffi::Buffer<ffi::F32> y,
                               ffi::ResultBuffer<ffi::PRED> out) {
    auto dx = x.dimensions();
    auto dy = y.dimensions();
    return status(evx_dominate_relation_u8(x.typed_data(), dx[0], y.typed_data(), dy[0], (int32_t)dx[1],
                                           reinterpret_cast<uint8_t*>(out->untyped_data()), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(EvxDominateRelation, RelationImpl,
                              ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>().Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>().Ret<ffi::Buffer<ffi::PRED>>());

// DTLZk.evaluate / MaF1.evaluate (which = 8)                       problems/numerical/dtlz.py, maf.py:95-108
static ffi::Error EvalDtlzImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> X, ffi::ResultBuffer<ffi::F32> F, int64_t which) {
    auto d = X.dimensions();
    const int32_t m = (int32_t)F->dimensions()[1];
    if (which == 8) return status(evx_eval_maf1_f32(X.typed_data(), d[0], (int32_t)d[1], m, F->typed_data(), stream));
    return status(evx_eval_dtlz_f32((int32_t)which, X.typed_data(), d[0], (int32_t)d[1], m, F->typed_data(), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(EvxEvalDtlz, EvalDtlzImpl,
                              ffi::Ffi::Bind().Ctx<ffi::PlatformStream<cudaStream_t>>().Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>().Attr<int64_t>("which"));
