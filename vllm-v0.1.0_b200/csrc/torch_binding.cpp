// pa_torch_ops: the reference-side stub in compiled form.
//
// A pybind11 module with the reference's operator signatures (csrc/attention.cpp:3-19,
// csrc/cache.cpp:6-47 of the reference: torch::Tensor& arguments, std::map block mappings,
// results in place) whose bodies only unpack pointers / sizes / the current stream and forward
// to the C ABI of libpaged_b200.so (include/paged_attention_b200.h).  No kernel code lives
// here.  It exists because a decode step calls these operators once per layer: argument
// unpacking in C++ costs ~2 us per call where the ctypes route (vllm-v0.1.0_b200/_cabi.py,
// same C ABI, kept for torch-free use and for tests without this module) costs ~30 us.
// Validation mirrors the reference's TORCH_CHECKs and adds the shape checks the reference
// leaves to luck; every failure surfaces as RuntimeError.
#include <torch/extension.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>

#include <map>
#include <vector>

#include "paged_attention_b200.h"

namespace {

int dtype_code(const torch::Tensor& t) {
  switch (t.scalar_type()) {
    case at::ScalarType::Half: return PA_F16;
    case at::ScalarType::BFloat16: return PA_BF16;
    case at::ScalarType::Float: return PA_F32;
    default: TORCH_CHECK(false, "Unsupported data type: ", t.scalar_type());
  }
  return -1;
}

void check(int rc) { TORCH_CHECK(rc == PA_OK, "libpaged_b200: ", pa_last_error(), " (status ", rc, ")"); }

void require_cuda(const char* name, const torch::Tensor& t, const c10::Device& dev) {
  TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor (there is no CPU path)");
  TORCH_CHECK(t.device() == dev, name, " is on ", t.device(), ", expected ", dev);
}

void* stream_of(const c10::Device& dev) { return at::cuda::getCurrentCUDAStream(dev.index()).stream(); }

void single_query_cached_kv_attention(torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache,
                                      torch::Tensor& value_cache, float scale, torch::Tensor& block_tables,
                                      torch::Tensor& context_lens, int block_size, int max_context_len,
                                      int variant, int num_warps, int num_stages, int num_splits) {
  const c10::Device dev = query.device();
  require_cuda("query", query, dev);
  require_cuda("out", out, dev);
  require_cuda("key_cache", key_cache, dev);
  require_cuda("value_cache", value_cache, dev);
  require_cuda("block_tables", block_tables, dev);
  require_cuda("context_lens", context_lens, dev);
  const int code = dtype_code(query);
  TORCH_CHECK(out.scalar_type() == query.scalar_type() && key_cache.scalar_type() == query.scalar_type() &&
                  value_cache.scalar_type() == query.scalar_type(),
              "out, query, key_cache and value_cache must share one dtype");
  // the reference's data_ptr<int>() throws on anything else (attention_kernels.cu:358-359)
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int && context_lens.scalar_type() == at::ScalarType::Int,
              "block_tables and context_lens must be int32");
  TORCH_CHECK(query.dim() == 3 && out.dim() == 3 && key_cache.dim() == 5 && value_cache.dim() == 4 &&
                  block_tables.dim() == 2,
              "unexpected tensor rank");
  const int64_t num_seqs = query.size(0), num_heads = query.size(1), head_size = query.size(2);
  TORCH_CHECK(out.size(0) == num_seqs && out.size(1) == num_heads && out.size(2) == head_size && out.is_contiguous(),
              "out must be a contiguous [num_seqs, num_heads, head_size] tensor");
  TORCH_CHECK(query.stride(2) == 1 && query.stride(1) == head_size,
              "query must be contiguous in its last two dimensions");
  TORCH_CHECK(key_cache.is_contiguous() && value_cache.is_contiguous() && block_tables.is_contiguous() &&
                  context_lens.is_contiguous(),
              "caches, block_tables and context_lens must be contiguous");
  TORCH_CHECK(value_cache.size(1) == num_heads && value_cache.size(2) == head_size && value_cache.size(3) == block_size,
              "value_cache shape does not match [num_blocks, num_heads, head_size, block_size]");
  TORCH_CHECK(block_tables.size(0) == num_seqs && context_lens.numel() == num_seqs,
              "block_tables / context_lens row count does not match query");
  const c10::cuda::CUDAGuard guard(dev);
  pa_decode_tuning tuning = {variant, num_warps, num_stages, num_splits};
  check(pa_decode_attention_ex(out.data_ptr(), query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), scale,
                               block_tables.data_ptr<int>(), context_lens.data_ptr<int>(),
                               static_cast<int32_t>(num_seqs), static_cast<int32_t>(num_heads),
                               static_cast<int32_t>(head_size), block_size, static_cast<int32_t>(block_tables.size(1)),
                               max_context_len, query.stride(0), code, &tuning, stream_of(dev)));
}

struct KvRows {
  c10::Device dev;
  int code;
  int32_t num_tokens, num_heads, head_size, block_size;
};

KvRows kv_rows(const char* op, const torch::Tensor& key, const torch::Tensor& value, const torch::Tensor& key_cache,
               const torch::Tensor& value_cache, const torch::Tensor& slot_mapping) {
  const c10::Device dev = key_cache.device();
  require_cuda("key", key, dev);
  require_cuda("value", value, dev);
  require_cuda("key_cache", key_cache, dev);
  require_cuda("value_cache", value_cache, dev);
  require_cuda("slot_mapping", slot_mapping, dev);
  const int code = dtype_code(key);
  TORCH_CHECK(value.scalar_type() == key.scalar_type() && key_cache.scalar_type() == key.scalar_type() &&
                  value_cache.scalar_type() == key.scalar_type(),
              op, ": key, value and both caches must share one dtype");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Int, op, ": slot_mapping must be int32");
  TORCH_CHECK(key.dim() == 3 && value.sizes() == key.sizes() && key_cache.dim() == 5 && value_cache.dim() == 4,
              op, ": unexpected tensor rank or shape");
  const int64_t n = key.size(0), h = key.size(1), d = key.size(2);
  TORCH_CHECK(key.stride(2) == 1 && key.stride(1) == d && value.stride(2) == 1 && value.stride(1) == d,
              op, ": key and value must be contiguous in their last two dimensions");
  TORCH_CHECK(key_cache.is_contiguous() && value_cache.is_contiguous() && slot_mapping.is_contiguous(),
              op, ": caches and slot_mapping must be contiguous");
  const int64_t block_size = key_cache.size(3), x = key_cache.size(4);   // as the reference: cache_kernels.cu:196-197
  TORCH_CHECK(x == 16 / static_cast<int64_t>(key.element_size()) && key_cache.size(1) == h && key_cache.size(2) * x == d,
              op, ": key_cache shape does not match [blocks, heads, head_size/x, block_size, x]");
  TORCH_CHECK(value_cache.size(1) == h && value_cache.size(2) == d && value_cache.size(3) == block_size,
              op, ": value_cache shape does not match [blocks, heads, head_size, block_size]");
  TORCH_CHECK(slot_mapping.numel() >= n, op, ": slot_mapping is shorter than the token count");
  return KvRows{dev, code, static_cast<int32_t>(n), static_cast<int32_t>(h), static_cast<int32_t>(d),
                static_cast<int32_t>(block_size)};
}

void reshape_and_cache(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache, torch::Tensor& value_cache,
                       torch::Tensor& slot_mapping, int64_t slot_pattern) {
  const KvRows r = kv_rows("reshape_and_cache", key, value, key_cache, value_cache, slot_mapping);
  const c10::cuda::CUDAGuard guard(r.dev);
  check(pa_reshape_and_cache_ex(key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                                slot_mapping.data_ptr<int>(), r.num_tokens, r.num_heads, r.head_size, r.block_size,
                                key.stride(0), value.stride(0), r.code, static_cast<int32_t>(slot_pattern),
                                stream_of(r.dev)));
}

void gather_cached_kv(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache, torch::Tensor& value_cache,
                      torch::Tensor& slot_mapping) {
  const KvRows r = kv_rows("gather_cached_kv", key, value, key_cache, value_cache, slot_mapping);
  const c10::cuda::CUDAGuard guard(r.dev);
  check(pa_gather_cached_kv(key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                            slot_mapping.data_ptr<int>(), r.num_tokens, r.num_heads, r.head_size, r.block_size,
                            key.stride(0), value.stride(0), r.code, stream_of(r.dev)));
}

void copy_blocks(std::vector<torch::Tensor>& key_caches, std::vector<torch::Tensor>& value_caches,
                 const std::map<int64_t, std::vector<int64_t>>& block_mapping) {
  const size_t num_layers = key_caches.size();
  TORCH_CHECK(num_layers == value_caches.size(), "copy_blocks: key_caches and value_caches differ in length");
  if (num_layers == 0) return;                                     // reference: cache_kernels.cu:88-90
  TORCH_CHECK(key_caches[0].is_cuda(), "copy_blocks: caches must be CUDA tensors");
  const c10::Device dev = key_caches[0].device();
  const int64_t block_bytes = key_caches[0][0].numel() * key_caches[0].element_size();
  std::vector<void*> kptrs(num_layers), vptrs(num_layers);
  for (size_t l = 0; l < num_layers; ++l) {
    const torch::Tensor& kc = key_caches[l];
    const torch::Tensor& vc = value_caches[l];
    TORCH_CHECK(kc.device() == dev && vc.device() == dev && kc.is_contiguous() && vc.is_contiguous(),
                "copy_blocks: every cache must be contiguous and on one GPU");
    TORCH_CHECK(kc.size(0) > 0 && vc.size(0) > 0 &&
                    kc.numel() / kc.size(0) * static_cast<int64_t>(kc.element_size()) == block_bytes &&
                    vc.numel() / vc.size(0) * static_cast<int64_t>(vc.element_size()) == block_bytes,
                "copy_blocks: caches differ in block size");
    kptrs[l] = kc.data_ptr();
    vptrs[l] = vc.data_ptr();
  }
  std::vector<int64_t> pairs;                                       // std::map order, as cache_kernels.cu:103-112
  for (const auto& kv : block_mapping)
    for (int64_t dst : kv.second) {
      pairs.push_back(kv.first);
      pairs.push_back(dst);
    }
  if (pairs.empty()) return;
  const c10::cuda::CUDAGuard guard(dev);
  check(pa_copy_blocks(kptrs.data(), vptrs.data(), static_cast<int32_t>(num_layers), pairs.data(),
                       static_cast<int64_t>(pairs.size() / 2), block_bytes, stream_of(dev)));
}

void swap_blocks(torch::Tensor& src, torch::Tensor& dst, const std::map<int64_t, int64_t>& block_mapping, int engine) {
  int kind;
  c10::Device dev = src.device();
  if (src.is_cuda() && dst.is_cuda()) {
    TORCH_CHECK(src.device() == dst.device(), "src and dst must be on the same GPU");   // cache_kernels.cu:17-19
    kind = PA_COPY_D2D;
  } else if (src.is_cuda() && dst.is_cpu()) {
    kind = PA_COPY_D2H;
  } else if (src.is_cpu() && dst.is_cuda()) {
    kind = PA_COPY_H2D;
    dev = dst.device();
  } else {
    TORCH_CHECK(false, "Invalid device combination");                                   // cache_kernels.cu:26
  }
  TORCH_CHECK(src.is_contiguous() && dst.is_contiguous(), "swap_blocks: caches must be contiguous");
  if (block_mapping.empty()) return;
  TORCH_CHECK(src.size(0) > 0 && dst.size(0) > 0, "swap_blocks: empty cache");
  const int64_t block_bytes = src.element_size() * (src.numel() / src.size(0));         // cache_kernels.cu:32
  TORCH_CHECK(static_cast<int64_t>(dst.element_size()) * (dst.numel() / dst.size(0)) == block_bytes,
              "swap_blocks: src and dst differ in block size");
  std::vector<int64_t> srcs, dsts;
  srcs.reserve(block_mapping.size());
  dsts.reserve(block_mapping.size());
  for (const auto& kv : block_mapping) {
    srcs.push_back(kv.first);
    dsts.push_back(kv.second);
  }
  const c10::cuda::CUDAGuard guard(dev);
  check(pa_swap_blocks_ex(src.data_ptr(), dst.data_ptr(), block_bytes, srcs.data(), dsts.data(),
                          static_cast<int64_t>(srcs.size()), kind, engine, stream_of(dev)));
}

struct RopeArgs {
  c10::Device dev;
  int code;
  int32_t num_tokens, num_heads, rot_dim;
};

RopeArgs rope_args(const char* op, const torch::Tensor& positions, const torch::Tensor& query, const torch::Tensor& key,
                   int head_size, const torch::Tensor& cos_sin_cache) {
  const c10::Device dev = query.device();
  require_cuda("positions", positions, dev);
  require_cuda("query", query, dev);
  require_cuda("key", key, dev);
  require_cuda("cos_sin_cache", cos_sin_cache, dev);
  const int code = dtype_code(query);
  TORCH_CHECK(key.scalar_type() == query.scalar_type() && cos_sin_cache.scalar_type() == query.scalar_type(),
              op, ": query, key and cos_sin_cache must share one dtype");
  TORCH_CHECK(positions.scalar_type() == at::ScalarType::Long, op, ": positions must be int64");
  TORCH_CHECK(query.dim() == 2 && key.dim() == 2 && cos_sin_cache.dim() == 2 && query.sizes() == key.sizes(),
              op, ": query and key must be [num_tokens, num_heads * head_size]");
  TORCH_CHECK(query.stride(1) == 1 && key.stride(1) == 1 && cos_sin_cache.is_contiguous() && positions.is_contiguous(),
              op, ": rows must be contiguous");
  TORCH_CHECK(head_size > 0 && query.size(1) % head_size == 0, op, ": hidden size ", query.size(1),
              " is not a multiple of head_size ", head_size);
  TORCH_CHECK(positions.numel() >= query.size(0), op, ": positions is shorter than the token count");
  return RopeArgs{dev, code, static_cast<int32_t>(query.size(0)), static_cast<int32_t>(query.size(1) / head_size),
                  static_cast<int32_t>(cos_sin_cache.size(1))};
}

void rotary_embedding_neox(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int head_size,
                           torch::Tensor& cos_sin_cache) {
  const RopeArgs r = rope_args("rotary_embedding_neox", positions, query, key, head_size, cos_sin_cache);
  const c10::cuda::CUDAGuard guard(r.dev);
  check(pa_rotary_embedding_neox(positions.data_ptr<int64_t>(), query.data_ptr(), key.data_ptr(),
                                 cos_sin_cache.data_ptr(), r.num_tokens, r.num_heads, head_size, r.rot_dim,
                                 query.stride(0), key.stride(0), r.code, stream_of(r.dev)));
}

void rotary_embedding_neox_reshape_and_cache(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key,
                                             torch::Tensor& value, torch::Tensor& key_cache,
                                             torch::Tensor& value_cache, torch::Tensor& slot_mapping, int head_size,
                                             torch::Tensor& cos_sin_cache) {
  const char* op = "rotary_embedding_neox_reshape_and_cache";
  const RopeArgs r = rope_args(op, positions, query, key, head_size, cos_sin_cache);
  require_cuda("value", value, r.dev);
  require_cuda("key_cache", key_cache, r.dev);
  require_cuda("value_cache", value_cache, r.dev);
  require_cuda("slot_mapping", slot_mapping, r.dev);
  TORCH_CHECK(value.scalar_type() == query.scalar_type() && key_cache.scalar_type() == query.scalar_type() &&
                  value_cache.scalar_type() == query.scalar_type(),
              op, ": value and both caches must have the dtype of query");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Int && slot_mapping.is_contiguous(),
              op, ": slot_mapping must be a contiguous int32 tensor");
  TORCH_CHECK(value.dim() == 2 && value.sizes() == key.sizes() && value.stride(1) == 1,
              op, ": value must be [num_tokens, num_heads * head_size] with contiguous rows");
  TORCH_CHECK(key_cache.dim() == 5 && value_cache.dim() == 4 && key_cache.is_contiguous() && value_cache.is_contiguous(),
              op, ": unexpected cache rank or layout");
  const int64_t block_size = key_cache.size(3), x = key_cache.size(4);
  TORCH_CHECK(x == 16 / static_cast<int64_t>(key.element_size()) && key_cache.size(1) == r.num_heads &&
                  key_cache.size(2) * x == head_size && value_cache.size(1) == r.num_heads &&
                  value_cache.size(2) == head_size && value_cache.size(3) == block_size,
              op, ": cache shapes do not match the key/value rows");
  TORCH_CHECK(slot_mapping.numel() <= r.num_tokens, op, ": slot_mapping is longer than the token batch");
  const c10::cuda::CUDAGuard guard(r.dev);
  check(pa_rope_reshape_and_cache(positions.data_ptr<int64_t>(), query.data_ptr(), key.data_ptr(), value.data_ptr(),
                                  key_cache.data_ptr(), value_cache.data_ptr(), slot_mapping.data_ptr<int>(),
                                  cos_sin_cache.data_ptr(), r.num_tokens, static_cast<int32_t>(slot_mapping.numel()),
                                  r.num_heads, head_size, static_cast<int32_t>(block_size), r.rot_dim,
                                  query.stride(0), key.stride(0), value.stride(0), r.code, stream_of(r.dev)));
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  namespace py = pybind11;
  // compiled against one version of the C ABI header; refuse to run against a library built from another
  if (pa_abi_version() != PA_ABI_VERSION)
    throw std::runtime_error("pa_torch_ops was compiled for ABI " + std::to_string(PA_ABI_VERSION) +
                             " but libpaged_b200.so reports " + std::to_string(pa_abi_version()) +
                             ": rebuild with python vllm-v0.1.0_b200/build.py");
  m.def("single_query_cached_kv_attention", &single_query_cached_kv_attention, py::arg("out"), py::arg("query"),
        py::arg("key_cache"), py::arg("value_cache"), py::arg("scale"), py::arg("block_tables"),
        py::arg("context_lens"), py::arg("block_size"), py::arg("max_context_len"), py::arg("variant") = 0,
        py::arg("num_warps") = 0, py::arg("num_stages") = 0, py::arg("num_splits") = 0,
        "Paged decode attention (replaces attention_ops.single_query_cached_kv_attention)");
  m.def("reshape_and_cache", &reshape_and_cache, py::arg("key"), py::arg("value"), py::arg("key_cache"),
        py::arg("value_cache"), py::arg("slot_mapping"), py::arg("slot_pattern") = 0,
        "Scatter new K/V rows into the paged caches");
  m.def("gather_cached_kv", &gather_cached_kv, "Gather K/V rows out of the paged caches");
  m.def("copy_blocks", &copy_blocks, "Copy-on-write of cache blocks in every layer");
  m.def("swap_blocks", &swap_blocks, py::arg("src"), py::arg("dst"), py::arg("block_mapping"), py::arg("engine") = 0,
        "Move cache blocks between GPU and host caches");
  m.def("rotary_embedding_neox", &rotary_embedding_neox, "GPT-NeoX rotary embedding in place on query and key");
  m.def("rotary_embedding_neox_reshape_and_cache", &rotary_embedding_neox_reshape_and_cache,
        "rotary_embedding_neox followed by reshape_and_cache, one launch");
  m.def("launch_count", []() { return pa_launch_count(); });
}
