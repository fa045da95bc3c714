// torch_ops.cpp -- the torch.library ops of the package, registered from C++ over the C ABI.
//
// Stand-in for the reference's popart plugin types (Op / GradOp / Opx / GradOpx / OpCreator,
// flash_attention_ipu/cpp/flash_attention_qkv_packed.cpp:516-676): op definition + shape rules
// (:571-577, :530), autodiff wiring (:536-546, :579-583) and lowering (:633-657) -- here: schema,
// CUDA implementation that fills fa_b200_problem / fa_b200_tensor and calls fa_b200_fwd / _bwd
// (include/fa_b200.h), and an autograd function.  Registered natively (not through Python's
// torch.library.custom_op) because that layer costs ~65 us of host time per forward+backward
// pair, which is what the small shapes (nanoGPT demo, GPT-2 small) are bound by.
//
// Autograd contract: forward returns (out, lse) and saves them; the reference's grad op
// receives only (dOut, qkv) and re-runs the forward (:357-359, :536-541) -- not replicated.
//
// Links against libflash_attention.so only through the C ABI; no kernels in this file.
#include <ATen/ATen.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/csrc/autograd/custom_function.h>
#include <torch/library.h>

#include <optional>
#include <tuple>
#include <vector>

#include "../../include/fa_b200.h"

namespace {

using at::Tensor;

int dtype_code(const Tensor &t) {
  switch (t.scalar_type()) {
    case at::kHalf: return FA_B200_F16;
    case at::kBFloat16: return FA_B200_BF16;
    case at::kFloat: return FA_B200_F32;
    default: TORCH_CHECK(false, "unsupported dtype ", t.scalar_type());
  }
  return -1;
}

void require_cuda(std::initializer_list<const Tensor *> ts) {
  for (const Tensor *t : ts)
    TORCH_CHECK(t->is_cuda(),
                "flash_attention_ipu_b200 runs on CUDA (sm_100a) tensors only; there is no CPU path");
}

// (B,H,L,D) view the kernels can address in place: unit stride on D, 16-byte aligned base, and for
// every other dim of size > 1 a positive, 16-byte aligned stride (the TMA path; an expanded dim has
// stride 0 and is materialised).  Size-1 dims never enter the address arithmetic, whatever
// stride torch reports for them: they are re-strided to the dense value so that the descriptor
// encoder sees legal numbers.  Anything else is made contiguous.
Tensor as_bhld(const Tensor &t) {
  const int64_t es = t.element_size();
  bool ok = t.dim() == 4 && t.stride(-1) == 1 && reinterpret_cast<uintptr_t>(t.data_ptr()) % 16 == 0;
  bool restride = false;
  for (int64_t i = 0; ok && i + 1 < t.dim(); ++i) {
    if (t.size(i) > 1) ok = t.stride(i) > 0 && (t.stride(i) * es) % 16 == 0;
    else if (t.stride(i) <= 0 || (t.stride(i) * es) % 16) restride = true;
  }
  if (!ok) return t.contiguous();
  if (!restride) return t;
  std::vector<int64_t> st(t.strides().begin(), t.strides().end());
  for (int64_t i = t.dim() - 2; i >= 0; --i)
    if (t.size(i) == 1) st[i] = st[i + 1] * t.size(i + 1);  // dense value; never used for addressing
  return t.as_strided(t.sizes(), st);
}

fa_b200_tensor desc(const Tensor &t) { return fa_b200_tensor{t.data_ptr(), t.stride(0), t.stride(1), t.stride(2)}; }

// q, k, v (and out / dout / lse when given) must describe ONE problem: same dtype, device, head
// dim, batch; v shaped like k; out / dout shaped like q; lse fp32 (B, H, Lq)
void check_operands(const Tensor &q, const Tensor &k, const Tensor &v, const Tensor *out, const Tensor *dout,
                    const Tensor *lse) {
  TORCH_CHECK(q.dim() == 4 && k.dim() == 4 && v.dim() == 4, "expected (B, H, L, D) operands");
  TORCH_CHECK(k.scalar_type() == q.scalar_type() && v.scalar_type() == q.scalar_type(),
              "q, k, v must share a dtype, got ", q.scalar_type(), " ", k.scalar_type(), " ", v.scalar_type());
  TORCH_CHECK(k.device() == q.device() && v.device() == q.device(), "q, k, v must be on one device");
  TORCH_CHECK(k.size(0) == q.size(0) && k.size(3) == q.size(3), "k of shape ", k.sizes(), " does not match q ", q.sizes());
  TORCH_CHECK(v.sizes() == k.sizes(), "v of shape ", v.sizes(), " does not match k ", k.sizes());
  for (const Tensor *t : {out, dout}) {
    if (!t) continue;
    TORCH_CHECK(t->sizes() == q.sizes(), "out / dout of shape ", t->sizes(), " does not match q ", q.sizes());
    TORCH_CHECK(t->scalar_type() == q.scalar_type() && t->device() == q.device(), "out / dout must match q's dtype and device");
  }
  if (lse) {
    TORCH_CHECK(lse->scalar_type() == at::kFloat && lse->device() == q.device(), "lse must be an fp32 tensor on q's device");
    TORCH_CHECK(lse->numel() == q.size(0) * q.size(1) * q.size(2), "lse has ", lse->numel(), " elements, expected B*H*Lq = ",
                q.size(0) * q.size(1) * q.size(2));
  }
}

// kv_lens: optional int32 CUDA tensor with one entry per (batch, query head) group -- the keys of that
// group that exist (key padding).  Returned contiguous so that the caller keeps it alive over the launch.
Tensor lens_or_empty(const std::optional<Tensor> &kv_lens, const Tensor &q) {
  if (!kv_lens.has_value() || !kv_lens->defined()) return Tensor();
  TORCH_CHECK(kv_lens->is_cuda() && kv_lens->device() == q.device(), "kv_lens must be on q's device");
  TORCH_CHECK(kv_lens->scalar_type() == at::kInt, "kv_lens must be int32");
  TORCH_CHECK(kv_lens->numel() == q.size(0) * q.size(1), "kv_lens has ", kv_lens->numel(),
              " entries, expected one per (batch, head) group = ", q.size(0) * q.size(1));
  return kv_lens->contiguous();
}

fa_b200_problem make_problem(const Tensor &q, const Tensor &k, bool causal, double scale, int64_t block_q,
                             int64_t block_kv, const Tensor &lens = Tensor(), double dropout_p = 0.0,
                             int64_t dropout_seed = 0) {
  TORCH_CHECK(q.dim() == 4 && k.dim() == 4, "expected (B, H, L, D) operands");
  const int64_t H = q.size(1), Hkv = k.size(1);
  TORCH_CHECK(Hkv > 0 && H % Hkv == 0, H, " query heads are not a multiple of ", Hkv, " key/value heads");
  fa_b200_problem p{};
  p.B = static_cast<int32_t>(q.size(0));
  p.H = static_cast<int32_t>(H);
  p.Lq = static_cast<int32_t>(q.size(2));
  p.Lk = static_cast<int32_t>(k.size(2));
  p.D = static_cast<int32_t>(q.size(3));
  p.dtype = dtype_code(q);
  p.causal = causal ? 1 : 0;
  p.softmax_scale = static_cast<float>(scale);
  p.block_q_hint = static_cast<int32_t>(block_q);
  p.block_kv_hint = static_cast<int32_t>(block_kv);
  p.q_pos0 = 0;
  p.kv_pos0 = 0;
  p.H_kv = Hkv == H ? 0 : static_cast<int32_t>(Hkv);
  p.kv_lens = lens.defined() ? lens.data_ptr<int32_t>() : nullptr;
  p.dropout_p = static_cast<float>(dropout_p);
  p.dropout_seed = static_cast<uint64_t>(dropout_seed);
  return p;
}

// Additive score bias (the float form of attn_mask): (B|1, H|1, Lq, Lk) in q's dtype.  Returned in a form the
// kernels address in place (unit stride on Lk, 16-byte aligned rows), so that the caller keeps it alive.
Tensor bias_or_empty(const std::optional<Tensor> &bias, const Tensor &q, const Tensor &k) {
  if (!bias.has_value() || !bias->defined()) return Tensor();
  const Tensor &b = *bias;
  TORCH_CHECK(b.is_cuda() && b.device() == q.device(), "bias must be on q's device");
  TORCH_CHECK(b.scalar_type() == q.scalar_type(), "bias must have q's dtype, got ", b.scalar_type());
  TORCH_CHECK(b.dim() == 4 && b.size(2) == q.size(2) && b.size(3) == k.size(2) && (b.size(0) == 1 || b.size(0) == q.size(0)) &&
                  (b.size(1) == 1 || b.size(1) == q.size(1)),
              "bias of shape ", b.sizes(), " is not (B|1, H|1, Lq, Lk) for q ", q.sizes(), " and k ", k.sizes());
  TORCH_CHECK(!b.requires_grad(), "no gradient is produced for the attention bias");
  // a stride of 0 (an expanded dim) is fine on every dim but the last: the kernels only multiply by it
  bool ok = b.stride(3) == 1 && reinterpret_cast<uintptr_t>(b.data_ptr()) % 16 == 0 &&
            (b.stride(2) == 0 || (b.stride(2) % 8 == 0 && b.stride(2) >= b.size(3)));
  for (int64_t i = 0; ok && i < 2; ++i) ok = b.size(i) == 1 || (b.stride(i) >= 0 && b.stride(i) % 8 == 0);
  return ok ? b : b.contiguous();
}

void set_bias(fa_b200_problem &p, const Tensor &b) {
  if (!b.defined()) return;
  p.bias = b.data_ptr();
  p.bias_stride_b = b.size(0) > 1 ? b.stride(0) : 0;
  p.bias_stride_h = b.size(1) > 1 ? b.stride(1) : 0;
  p.bias_stride_l = b.stride(2);
}

void check(int rc, const char *what) {
  TORCH_CHECK(rc == 0, what, " failed (status ", rc, "): ", fa_b200_last_error());
}

void run_fwd(const fa_b200_problem &p, const Tensor &q, const Tensor &k, const Tensor &v, const Tensor &o,
             const Tensor &lse) {
  const c10::cuda::CUDAGuard guard(q.device());
  const fa_b200_tensor tq = desc(q), tk = desc(k), tv = desc(v), to = desc(o);
  check(fa_b200_fwd(&p, &tq, &tk, &tv, &to, lse.data_ptr<float>(),
                    c10::cuda::getCurrentCUDAStream(q.device().index()).stream()),
        "fa_b200_fwd");
}

void run_bwd(const fa_b200_problem &p, const Tensor &q, const Tensor &k, const Tensor &v, const Tensor &o,
             const Tensor &dout, const Tensor &lse, const Tensor &dq, const Tensor &dk, const Tensor &dv) {
  const c10::cuda::CUDAGuard guard(q.device());
  const int64_t ws_bytes = static_cast<int64_t>(fa_b200_bwd_workspace_bytes(&p));
  const Tensor ws = at::empty({ws_bytes}, q.options().dtype(at::kByte));
  const fa_b200_tensor tq = desc(q), tk = desc(k), tv = desc(v), to = desc(o), tdo = desc(dout);
  const fa_b200_tensor tdq = desc(dq), tdk = desc(dk), tdv = desc(dv);
  // torch.use_deterministic_algorithms(True) asks for the ordered dQ accumulation for this call
  const bool det = at::globalContext().deterministicAlgorithms();
  const int prev = det ? fa_b200_set_deterministic(1) : 0;
  const int rc = fa_b200_bwd(&p, &tq, &tk, &tv, &to, &tdo, lse.data_ptr<float>(), &tdq, &tdk, &tdv, ws.data_ptr(),
                             static_cast<size_t>(ws_bytes), c10::cuda::getCurrentCUDAStream(q.device().index()).stream());
  if (det) fa_b200_set_deterministic(prev);
  check(rc, "fa_b200_bwd");
}

// ------------------------------------------------------------------ separate q, k, v (SDPA path)
std::tuple<Tensor, Tensor> attention_fwd(const Tensor &q_, const Tensor &k_, const Tensor &v_, double scale,
                                         bool causal, int64_t block_q, int64_t block_kv,
                                         const std::optional<Tensor> &kv_lens, double dropout_p, int64_t dropout_seed,
                                         const std::optional<Tensor> &bias_) {
  require_cuda({&q_, &k_, &v_});
  check_operands(q_, k_, v_, nullptr, nullptr, nullptr);
  const Tensor q = as_bhld(q_), k = as_bhld(k_), v = as_bhld(v_);
  Tensor out = at::empty_like(q);  // keeps q's (possibly (B,L,H,D)-transposed) dense layout
  if (out.stride(-1) != 1) out = at::empty(q.sizes(), q.options());
  const Tensor lse = at::empty({q.size(0), q.size(1), q.size(2)}, q.options().dtype(at::kFloat));
  const Tensor lens = lens_or_empty(kv_lens, q);
  const Tensor bias = bias_or_empty(bias_, q, k);
  fa_b200_problem prob = make_problem(q, k, causal, scale, block_q, block_kv, lens, dropout_p, dropout_seed);
  set_bias(prob, bias);
  run_fwd(prob, q, k, v, out, lse);
  return {out, lse};
}

std::tuple<Tensor, Tensor, Tensor> attention_bwd(const Tensor &dout_, const Tensor &q_, const Tensor &k_,
                                                 const Tensor &v_, const Tensor &out_, const Tensor &lse_,
                                                 double scale, bool causal, int64_t block_q, int64_t block_kv,
                                                 const std::optional<Tensor> &kv_lens, double dropout_p,
                                                 int64_t dropout_seed, const std::optional<Tensor> &bias_) {
  require_cuda({&dout_, &q_, &k_, &v_, &out_, &lse_});
  check_operands(q_, k_, v_, &out_, &dout_, &lse_);
  const Tensor lse = lse_.contiguous();  // the kernels index it as a dense [B*H][Lq] array
  const Tensor q = as_bhld(q_), k = as_bhld(k_), v = as_bhld(v_), out = as_bhld(out_), dout = as_bhld(dout_);
  const Tensor dq = at::empty_like(q), dk = at::empty_like(k), dv = at::empty_like(v);
  const Tensor lens = lens_or_empty(kv_lens, q);
  const Tensor bias = bias_or_empty(bias_, q, k);
  fa_b200_problem prob = make_problem(q, k, causal, scale, block_q, block_kv, lens, dropout_p, dropout_seed);
  set_bias(prob, bias);
  run_bwd(prob, q, k, v, out, dout, lse, dq, dk, dv);
  return {dq, dk, dv};
}

// ------------------------------------------------------------------ packed qkv (reference API)
// (3,G,L,D) is B=1, H=G with q/k/v = base + {0,1,2}*G*L*D: zero-copy views.
std::tuple<Tensor, Tensor> qkv_packed_fwd(const Tensor &qkv_, double scale, bool causal, int64_t block_q,
                                          int64_t block_kv, const std::optional<Tensor> &kv_lens, double dropout_p,
                                          int64_t dropout_seed) {
  require_cuda({&qkv_});
  TORCH_CHECK(qkv_.dim() == 4 && qkv_.size(0) == 3, "expected a packed (3, G, L, D) tensor");
  const Tensor qkv = qkv_.contiguous();
  const Tensor q = qkv[0].unsqueeze(0), k = qkv[1].unsqueeze(0), v = qkv[2].unsqueeze(0);
  const Tensor out = at::empty({qkv.size(1), qkv.size(2), qkv.size(3)}, qkv.options());
  const Tensor lse = at::empty({qkv.size(1), qkv.size(2)}, qkv.options().dtype(at::kFloat));
  const Tensor lens = lens_or_empty(kv_lens, q);
  run_fwd(make_problem(q, k, causal, scale, block_q, block_kv, lens, dropout_p, dropout_seed), q, k, v, out.unsqueeze(0), lse);
  return {out, lse};
}

Tensor qkv_packed_bwd(const Tensor &dout_, const Tensor &qkv_, const Tensor &out_, const Tensor &lse_,
                      double scale, bool causal, int64_t block_q, int64_t block_kv,
                      const std::optional<Tensor> &kv_lens, double dropout_p, int64_t dropout_seed) {
  require_cuda({&dout_, &qkv_, &out_, &lse_});
  TORCH_CHECK(qkv_.dim() == 4 && qkv_.size(0) == 3, "expected a packed (3, G, L, D) tensor");
  TORCH_CHECK(dout_.dim() == 3 && out_.dim() == 3 && dout_.sizes() == out_.sizes() && out_.size(0) == qkv_.size(1) &&
                  out_.size(1) == qkv_.size(2) && out_.size(2) == qkv_.size(3),
              "out / dout must be (G, L, D) tensors matching the packed qkv ", qkv_.sizes());
  TORCH_CHECK(dout_.scalar_type() == qkv_.scalar_type() && out_.scalar_type() == qkv_.scalar_type(), "dtype mismatch");
  TORCH_CHECK(lse_.scalar_type() == at::kFloat && lse_.numel() == qkv_.size(1) * qkv_.size(2), "lse must be fp32 (G, L)");
  const Tensor lse = lse_.contiguous();
  const Tensor qkv = qkv_.contiguous(), dout = dout_.contiguous(), out = out_.contiguous();
  const Tensor q = qkv[0].unsqueeze(0), k = qkv[1].unsqueeze(0), v = qkv[2].unsqueeze(0);
  const Tensor dqkv = at::empty_like(qkv);  // gradients land directly in the packed buffer
  const Tensor lens = lens_or_empty(kv_lens, q);
  run_bwd(make_problem(q, k, causal, scale, block_q, block_kv, lens, dropout_p, dropout_seed), q, k, v, out.unsqueeze(0), dout.unsqueeze(0), lse,
          dqkv[0].unsqueeze(0), dqkv[1].unsqueeze(0), dqkv[2].unsqueeze(0));
  return dqkv;
}

// ------------------------------------------------------------------ autograd
using torch::autograd::AutogradContext;
using torch::autograd::variable_list;

template <typename Sig>
auto typed_op(const char *name) {
  return c10::Dispatcher::singleton().findSchemaOrThrow(name, "").typed<Sig>();
}

struct AttentionFn : public torch::autograd::Function<AttentionFn> {
  static variable_list forward(AutogradContext *ctx, const Tensor &q, const Tensor &k, const Tensor &v,
                               double scale, bool causal, int64_t block_q, int64_t block_kv,
                               const std::optional<Tensor> &kv_lens, double dropout_p, int64_t dropout_seed,
                               const std::optional<Tensor> &bias) {
    at::AutoDispatchBelowADInplaceOrView below;
    static auto op = typed_op<std::tuple<Tensor, Tensor>(const Tensor &, const Tensor &, const Tensor &, double,
                                                         bool, int64_t, int64_t, const std::optional<Tensor> &, double,
                                                         int64_t, const std::optional<Tensor> &)>(
        "flash_attention_ipu::attention_fwd");
    auto [out, lse] = op.call(q, k, v, scale, causal, block_q, block_kv, kv_lens, dropout_p, dropout_seed, bias);
    ctx->saved_data["dropout_p"] = dropout_p;      // the backward regenerates the forward's mask from the seed
    ctx->saved_data["dropout_seed"] = dropout_seed;
    ctx->save_for_backward({q, k, v, out, lse, kv_lens.has_value() ? *kv_lens : Tensor(),
                            bias.has_value() ? *bias : Tensor()});
    ctx->saved_data["scale"] = scale;
    ctx->saved_data["causal"] = causal;
    ctx->saved_data["block_q"] = block_q;
    ctx->saved_data["block_kv"] = block_kv;
    ctx->mark_non_differentiable({lse});
    return {out, lse};
  }
  static variable_list backward(AutogradContext *ctx, variable_list grads) {
    static auto op = typed_op<std::tuple<Tensor, Tensor, Tensor>(
        const Tensor &, const Tensor &, const Tensor &, const Tensor &, const Tensor &, const Tensor &, double, bool,
        int64_t, int64_t, const std::optional<Tensor> &, double, int64_t, const std::optional<Tensor> &)>(
        "flash_attention_ipu::attention_bwd");
    const auto saved = ctx->get_saved_variables();
    const Tensor dout = grads[0].defined() ? grads[0] : at::zeros_like(saved[3]);
    const std::optional<Tensor> lens = saved[5].defined() ? std::optional<Tensor>(saved[5]) : std::nullopt;
    const std::optional<Tensor> bias = saved[6].defined() ? std::optional<Tensor>(saved[6]) : std::nullopt;
    auto [dq, dk, dv] = op.call(dout, saved[0], saved[1], saved[2], saved[3], saved[4],
                                ctx->saved_data["scale"].toDouble(), ctx->saved_data["causal"].toBool(),
                                ctx->saved_data["block_q"].toInt(), ctx->saved_data["block_kv"].toInt(), lens,
                                ctx->saved_data["dropout_p"].toDouble(), ctx->saved_data["dropout_seed"].toInt(), bias);
    return {dq, dk, dv, Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor()};
  }
};

struct PackedFn : public torch::autograd::Function<PackedFn> {
  static variable_list forward(AutogradContext *ctx, const Tensor &qkv, double scale, bool causal,
                               int64_t block_q, int64_t block_kv, const std::optional<Tensor> &kv_lens,
                               double dropout_p, int64_t dropout_seed) {
    at::AutoDispatchBelowADInplaceOrView below;
    static auto op = typed_op<std::tuple<Tensor, Tensor>(const Tensor &, double, bool, int64_t, int64_t,
                                                         const std::optional<Tensor> &, double, int64_t)>(
        "flash_attention_ipu::qkv_packed_fwd");
    auto [out, lse] = op.call(qkv, scale, causal, block_q, block_kv, kv_lens, dropout_p, dropout_seed);
    ctx->saved_data["dropout_p"] = dropout_p;
    ctx->saved_data["dropout_seed"] = dropout_seed;
    ctx->save_for_backward({qkv, out, lse, kv_lens.has_value() ? *kv_lens : Tensor()});
    ctx->saved_data["scale"] = scale;
    ctx->saved_data["causal"] = causal;
    ctx->saved_data["block_q"] = block_q;
    ctx->saved_data["block_kv"] = block_kv;
    ctx->mark_non_differentiable({lse});
    return {out, lse};
  }
  static variable_list backward(AutogradContext *ctx, variable_list grads) {
    static auto op = typed_op<Tensor(const Tensor &, const Tensor &, const Tensor &, const Tensor &, double, bool,
                                     int64_t, int64_t, const std::optional<Tensor> &, double, int64_t)>(
        "flash_attention_ipu::qkv_packed_bwd");
    const auto saved = ctx->get_saved_variables();
    const Tensor dout = grads[0].defined() ? grads[0] : at::zeros_like(saved[1]);
    const std::optional<Tensor> lens = saved[3].defined() ? std::optional<Tensor>(saved[3]) : std::nullopt;
    const Tensor dqkv = op.call(dout, saved[0], saved[1], saved[2], ctx->saved_data["scale"].toDouble(),
                                ctx->saved_data["causal"].toBool(), ctx->saved_data["block_q"].toInt(),
                                ctx->saved_data["block_kv"].toInt(), lens, ctx->saved_data["dropout_p"].toDouble(),
                                ctx->saved_data["dropout_seed"].toInt());
    return {dqkv, Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor()};
  }
};

std::tuple<Tensor, Tensor> attention_fwd_autograd(const Tensor &q, const Tensor &k, const Tensor &v, double scale,
                                                  bool causal, int64_t block_q, int64_t block_kv,
                                                  const std::optional<Tensor> &kv_lens, double dropout_p,
                                                  int64_t dropout_seed, const std::optional<Tensor> &bias) {
  auto r = AttentionFn::apply(q, k, v, scale, causal, block_q, block_kv, kv_lens, dropout_p, dropout_seed, bias);
  return {r[0], r[1]};
}

std::tuple<Tensor, Tensor> qkv_packed_fwd_autograd(const Tensor &qkv, double scale, bool causal, int64_t block_q,
                                                   int64_t block_kv, const std::optional<Tensor> &kv_lens,
                                                   double dropout_p, int64_t dropout_seed) {
  auto r = PackedFn::apply(qkv, scale, causal, block_q, block_kv, kv_lens, dropout_p, dropout_seed);
  return {r[0], r[1]};
}

}  // namespace

TORCH_LIBRARY(flash_attention_ipu, m) {
  // kv_lens (optional, int32, one entry per (batch, head) group): key padding -- only the first kv_lens[g]
  // keys of group g exist.  The reference refuses every attn_mask (_impl/scaled_dot_product_attention.py:21-24).
  // dropout_p / dropout_seed: dropout on the attention probabilities inside the tcgen05 kernels (the reference
  // refuses dropout_p > 0, :25-28); the mask is a function of the seed, so the backward op takes the same pair.
  // bias (attention_fwd / attention_bwd): additive score bias (B|1, H|1, Lq, Lk) in q's dtype, the float form of
  // attn_mask; no gradient is produced for it.
  m.def("attention_fwd(Tensor q, Tensor k, Tensor v, float softmax_scale, bool causal, int block_q, "
        "int block_kv, Tensor? kv_lens=None, float dropout_p=0.0, int dropout_seed=0, Tensor? bias=None) -> "
        "(Tensor, Tensor)");
  m.def("attention_bwd(Tensor dout, Tensor q, Tensor k, Tensor v, Tensor out, Tensor lse, float softmax_scale, "
        "bool causal, int block_q, int block_kv, Tensor? kv_lens=None, float dropout_p=0.0, int dropout_seed=0, "
        "Tensor? bias=None) -> (Tensor, Tensor, Tensor)");
  m.def("qkv_packed_fwd(Tensor qkv, float softmax_scale, bool causal, int block_q, int block_kv, "
        "Tensor? kv_lens=None, float dropout_p=0.0, int dropout_seed=0) -> (Tensor, Tensor)");
  m.def("qkv_packed_bwd(Tensor dout, Tensor qkv, Tensor out, Tensor lse, float softmax_scale, bool causal, "
        "int block_q, int block_kv, Tensor? kv_lens=None, float dropout_p=0.0, int dropout_seed=0) -> Tensor");
}

// The implementations refuse non-CUDA tensors themselves ("there is no CPU path"), so they are
// registered for every backend: a CPU tensor gets that message, not a dispatcher error.
TORCH_LIBRARY_IMPL(flash_attention_ipu, CompositeExplicitAutograd, m) {
  m.impl("attention_fwd", &attention_fwd);
  m.impl("attention_bwd", &attention_bwd);
  m.impl("qkv_packed_fwd", &qkv_packed_fwd);
  m.impl("qkv_packed_bwd", &qkv_packed_bwd);
}

TORCH_LIBRARY_IMPL(flash_attention_ipu, Autograd, m) {
  m.impl("attention_fwd", &attention_fwd_autograd);
  m.impl("qkv_packed_fwd", &qkv_packed_fwd_autograd);
}
