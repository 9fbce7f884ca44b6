// The host side of `experiment.compute_vtrace_loss` / `compute_learner_loss`
// in C++: argument checks, output allocation, the call into the C ABI of
// libvtrace_b200.so and the autograd node that hands out the gradients the
// forward launch already produced.
//
// This file holds no arithmetic and no kernel: it is the PyTorch plumbing of
// scalable_agent_b200/_ops.py (FusedLoss / FusedLearnerLoss) moved out of
// Python, because at the shapes of the reference's own learner batches
// (experiment.py:64-65: T=100, B=32) the Python version of that plumbing costs
// ten times what the kernel does.  The C ABI entry points are handed over as
// addresses by scalable_agent_b200/_ext.py (taken from the ctypes handle of the
// one loaded library), so there is a single copy of the library in the
// process and no link-time dependency.
//
// Replaces the same reference lines as the Python layer: experiment.py:362-407
// (loss block of build_learner) and :415 (backward).
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/vtrace_b200.h"

namespace {

using at::Tensor;
using torch::autograd::AutogradContext;
using torch::autograd::variable_list;

struct Abi {
  int (*fused)(const vtr_fused_params*, void*, size_t, void*) = nullptr;
  size_t (*workspace_bytes)(int32_t, int32_t, int32_t) = nullptr;
  int (*scale_inplace)(float*, const float*, int64_t, void*) = nullptr;
  const char* (*last_error)() = nullptr;
  int (*importance_weights)(const float*, const float*, const float*, const float*, const float*, int32_t, int32_t,
                            int32_t, uint32_t, float, float, uint32_t, float*, float*, void*) = nullptr;
} g_abi;

void bind(uintptr_t fused, uintptr_t workspace_bytes, uintptr_t scale_inplace, uintptr_t last_error) {
  g_abi.fused = reinterpret_cast<decltype(g_abi.fused)>(fused);
  g_abi.workspace_bytes = reinterpret_cast<decltype(g_abi.workspace_bytes)>(workspace_bytes);
  g_abi.scale_inplace = reinterpret_cast<decltype(g_abi.scale_inplace)>(scale_inplace);
  g_abi.last_error = reinterpret_cast<decltype(g_abi.last_error)>(last_error);
}

void bind_importance_weights(uintptr_t fn) {
  g_abi.importance_weights = reinterpret_cast<decltype(g_abi.importance_weights)>(fn);
}

// The C return convention as exceptions: negative = argument error
// (ValueError in Python), positive = a failed launch (RuntimeError).
void check(int rc, const char* what) {
  if (rc == VTR_OK) return;
  const std::string msg = std::string(what) + " failed (" + std::to_string(rc) + "): " + g_abi.last_error();
  TORCH_CHECK_VALUE(rc > 0, msg);
  TORCH_CHECK(false, msg);
}

void want(const Tensor& t, const char* name, at::ScalarType dtype, int64_t rank, const at::Device& dev) {
  TORCH_CHECK(t.is_cuda() && t.device() == dev, "b200vtrace ops need all tensors on one CUDA device (", name, ")");
  TORCH_CHECK(t.scalar_type() == dtype, name, " must be ", dtype);
  TORCH_CHECK_VALUE(t.dim() == rank, name, " must have rank ", rank, ", got shape ", t.sizes());
  TORCH_CHECK(t.is_contiguous(), "b200vtrace ops need contiguous tensors (", name, ")");
}

struct Options {
  c10::optional<double> clip_rho, clip_pg_rho;
  double baseline_cost, entropy_cost, discounting;
  int64_t reward_clip_mode, path;
  bool want_aux, pad_grads;
};

// One launch of vtr_from_logits_fused_f32 over already validated tensors.
// Returns {total, losses[4], vs, pg, log_rhos, b_lp, t_lp, grads} where grads
// is one block [d_target_logits | d_values] (empty without want_grads), each
// with Tg = T (+1 with pad_grads, last step zero) time steps.
std::vector<Tensor> launch(const Tensor& behaviour, const Tensor& target, const Tensor& actions,
                           const c10::optional<Tensor>& discounts, const Tensor& rewards,
                           const Tensor& values, const Tensor& bootstrap, const c10::optional<Tensor>& done,
                           const Options& o, bool want_grads) {
  TORCH_CHECK(g_abi.fused != nullptr, "vtrace_torch_ext: the C ABI was not bound");
  const at::Device dev = target.device();
  want(target, "target_policy_logits", at::kFloat, 3, dev);
  const int64_t T = target.size(0), B = target.size(1), A = target.size(2);
  want(behaviour, "behaviour_policy_logits", at::kFloat, 3, dev);
  TORCH_CHECK_VALUE(behaviour.sizes() == target.sizes(), "behaviour and target logits differ in shape");
  want(actions, "actions", at::kInt, 2, dev);
  want(rewards, "rewards", at::kFloat, 2, dev);
  want(values, "values", at::kFloat, 2, dev);
  want(bootstrap, "bootstrap_value", at::kFloat, 1, dev);
  const bool has_done = done.has_value() && done->defined();
  if (has_done) {
    want(*done, "done", at::kByte, 2, dev);
    TORCH_CHECK_VALUE(done->size(0) == T && done->size(1) == B, "done must have shape [T, B]");
  } else {
    TORCH_CHECK_VALUE(discounts.has_value() && discounts->defined(), "one of discounts / done must be given");
    want(*discounts, "discounts", at::kFloat, 2, dev);
    TORCH_CHECK_VALUE(discounts->size(0) == T && discounts->size(1) == B, "discounts must have shape [T, B]");
  }
  for (const Tensor* t : {&actions, &rewards, &values})
    TORCH_CHECK_VALUE(t->size(0) == T && t->size(1) == B, "actions, rewards and values must have shape [T, B] = [",
                      T, ", ", B, "]");
  TORCH_CHECK_VALUE(bootstrap.size(0) == B, "bootstrap_value must have shape [B]");
  TORCH_CHECK_VALUE(T > 0 && B > 0 && A > 0 && T * B < (int64_t(1) << 31), "bad shape");

  c10::cuda::CUDAGuard guard(dev);
  const auto f32 = target.options();
  // every [T, B] output in one block: vs, pg_advantages and, if asked for, the three aux fields
  const int64_t tb = T * B, fields = o.want_aux ? 5 : 2;
  Tensor tb_block = at::empty({fields, T, B}, f32);
  Tensor losses = at::empty({4}, f32);
  Tensor total = at::empty({}, f32);  // the differentiable output must not alias the constant ones
  const int64_t Tg = o.pad_grads ? T + 1 : T;
  Tensor grads = at::empty({want_grads ? Tg * B * (A + 1) : 0}, f32);

  vtr_fused_params p = {};
  p.T = static_cast<int32_t>(T), p.B = static_cast<int32_t>(B), p.A = static_cast<int32_t>(A);
  p.flags = (o.clip_rho ? 0u : VTR_FLAG_NO_CLIP_RHO) | (o.clip_pg_rho ? 0u : VTR_FLAG_NO_CLIP_PG_RHO);
  p.behaviour_policy_logits = behaviour.data_ptr<float>();
  p.target_policy_logits = target.data_ptr<float>();
  p.actions = actions.data_ptr<int32_t>();
  p.discounts = has_done ? nullptr : discounts->data_ptr<float>();
  p.rewards = rewards.data_ptr<float>();
  p.values = values.data_ptr<float>();
  p.bootstrap_value = bootstrap.data_ptr<float>();
  p.done = has_done ? done->data_ptr<uint8_t>() : nullptr;
  p.clip_rho_threshold = o.clip_rho ? static_cast<float>(*o.clip_rho) : 0.0f;
  p.clip_pg_rho_threshold = o.clip_pg_rho ? static_cast<float>(*o.clip_pg_rho) : 0.0f;
  p.baseline_cost = static_cast<float>(o.baseline_cost);
  p.entropy_cost = static_cast<float>(o.entropy_cost);
  p.discounting = static_cast<float>(o.discounting);
  p.reward_clip_mode = static_cast<int32_t>(o.reward_clip_mode);
  p.path = static_cast<int32_t>(o.path);
  float* out = tb_block.data_ptr<float>();
  p.vs = out, p.pg_advantages = out + tb;
  if (o.want_aux) p.log_rhos = out + 2 * tb, p.behaviour_action_log_probs = out + 3 * tb, p.target_action_log_probs = out + 4 * tb;
  if (want_grads) {
    float* g = grads.data_ptr<float>();
    p.d_target_logits = g;            // first: keeps the 16-byte alignment the bulk stores want
    p.d_values = g + Tg * B * A;
    if (o.pad_grads) {                // the step that only feeds the bootstrap value gets no gradient
      grads.narrow(0, T * B * A, B * A).zero_();
      grads.narrow(0, Tg * B * A + T * B, B).zero_();
    }
  }
  p.losses = losses.data_ptr<float>();
  p.total_loss = total.data_ptr<float>();
  const size_t ws_bytes = g_abi.workspace_bytes(p.T, p.B, p.A);
  Tensor workspace = at::empty({static_cast<int64_t>(ws_bytes / 8)}, f32.dtype(at::kDouble));
  cudaStream_t stream = at::cuda::getCurrentCUDAStream(dev.index()).stream();
  check(g_abi.fused(&p, workspace.data_ptr<double>(), ws_bytes, stream), "vtr_from_logits_fused_f32");

  // {total, losses, vs, pg_advantages, [log_rhos, b_lp, t_lp,] grads}
  std::vector<Tensor> r;
  r.reserve(8);
  r.push_back(total);
  r.push_back(losses);
  for (int64_t k = 0; k < fields; ++k) r.push_back(tb_block.select(0, k));
  r.push_back(grads);
  return r;
}

// Splits launch()'s result into the autograd outputs (all but the first
// constant) and the stashed gradient block.
variable_list outputs_of(AutogradContext* ctx, std::vector<Tensor>& r) {
  // Only the total ever receives a gradient.  Without this the engine would
  // hand backward() a freshly zero-filled tensor for each of the other
  // outputs - five [T, B] fills per step.
  ctx->set_materialize_grads(false);
  ctx->saved_data["grads"] = r.back();
  r.pop_back();
  ctx->mark_non_differentiable(variable_list(r.begin() + 1, r.end()));
  return r;
}

// Gradient views into the block of launch().
Tensor d_logits_of(const Tensor& grads, int64_t Tg, int64_t B, int64_t A) {
  return grads.narrow(0, 0, Tg * B * A).view({Tg, B, A});
}
Tensor d_values_of(const Tensor& grads, int64_t Tg, int64_t B, int64_t A) {
  return grads.narrow(0, Tg * B * A, Tg * B).view({Tg, B});
}

// Applies the upstream gradient of the total to the stashed gradients: one
// launch over the whole block, a no-op on the device when it is exactly 1.
void scale_block(const Tensor& grads, const Tensor& g_total) {
  if (grads.numel() == 0) return;
  Tensor g = g_total;
  if (g.scalar_type() != at::kFloat || !g.is_contiguous() || g.device() != grads.device())
    g = g.to(grads.device(), at::kFloat).contiguous();
  c10::cuda::CUDAGuard guard(grads.device());
  cudaStream_t stream = at::cuda::getCurrentCUDAStream(grads.device().index()).stream();
  check(g_abi.scale_inplace(grads.data_ptr<float>(), g.data_ptr<float>(), grads.numel(), stream),
        "vtr_scale_inplace_f32");
}

// experiment.py:389-407 + :415 on [T, ...] tensors.
class FusedLossFn : public torch::autograd::Function<FusedLossFn> {
 public:
  static variable_list forward(AutogradContext* ctx, const Tensor& behaviour, const Tensor& target,
                               const Tensor& actions, const c10::optional<Tensor>& discounts,
                               const Tensor& rewards, const Tensor& values, const Tensor& bootstrap,
                               const c10::optional<Tensor>& done, c10::optional<double> clip_rho,
                               c10::optional<double> clip_pg_rho, double baseline_cost, double entropy_cost,
                               double discounting, int64_t reward_clip_mode, bool want_aux, int64_t path) {
    const bool want_grads = target.requires_grad() || values.requires_grad();
    const Options o = {clip_rho, clip_pg_rho, baseline_cost, entropy_cost, discounting, reward_clip_mode, path,
                       want_aux, false};
    std::vector<Tensor> r = launch(behaviour, target, actions, discounts, rewards, values, bootstrap, done, o,
                                   want_grads);
    ctx->saved_data["dims"] = std::vector<int64_t>{target.size(0), target.size(1), target.size(2)};
    ctx->saved_data["need"] = std::vector<bool>{target.requires_grad(), values.requires_grad()};
    ctx->saved_data["ran"] = false;
    return outputs_of(ctx, r);
  }

  static variable_list backward(AutogradContext* ctx, variable_list grad_outputs) {
    TORCH_CHECK(!ctx->saved_data["ran"].toBool(),
                "FusedLoss hands out its stashed gradients in place; backward can run only once per forward");
    ctx->saved_data["ran"] = true;
    const Tensor grads = ctx->saved_data["grads"].toTensor();
    const auto dims = ctx->saved_data["dims"].toIntVector();
    const auto need = ctx->saved_data["need"].toBoolList();
    variable_list out(16);
    if (grads.numel() == 0 || !grad_outputs[0].defined()) return out;
    scale_block(grads, grad_outputs[0]);
    if (need[0]) out[1] = d_logits_of(grads, dims[0], dims[1], dims[2]);
    if (need[1]) out[5] = d_values_of(grads, dims[0], dims[1], dims[2]);
    return out;
  }
};

// experiment.py:362-407 on the learner's own [T+1, ...] tensors: the shifts
// of :366-375 are views, the gradients come back [T+1, ...] with a zero last
// step.
class FusedLearnerLossFn : public torch::autograd::Function<FusedLearnerLossFn> {
 public:
  static variable_list forward(AutogradContext* ctx, const Tensor& learner_logits, const Tensor& learner_baseline,
                               const Tensor& behaviour_logits, const Tensor& actions, const Tensor& rewards,
                               const Tensor& done, c10::optional<double> clip_rho,
                               c10::optional<double> clip_pg_rho, double baseline_cost, double entropy_cost,
                               double discounting, int64_t reward_clip_mode, bool want_aux, int64_t path) {
    TORCH_CHECK_VALUE(learner_logits.dim() == 3, "learner_policy_logits must have rank 3, got shape ",
                      learner_logits.sizes());
    const int64_t T = learner_logits.size(0) - 1, B = learner_logits.size(1), A = learner_logits.size(2);
    TORCH_CHECK_VALUE(T >= 1, "need at least 2 time steps (T+1), got ", T + 1);
    for (const Tensor* t : {&learner_baseline, &actions, &rewards, &done})
      TORCH_CHECK_VALUE(t->dim() == 2 && t->size(0) == T + 1 && t->size(1) == B,
                        "learner_baseline, actions, rewards and done must have shape [T+1, B]");
    TORCH_CHECK_VALUE(behaviour_logits.sizes() == learner_logits.sizes(),
                      "behaviour and learner logits differ in shape");
    const bool want_grads = learner_logits.requires_grad() || learner_baseline.requires_grad();
    const Options o = {clip_rho, clip_pg_rho, baseline_cost, entropy_cost, discounting, reward_clip_mode, path,
                       want_aux, true};
    std::vector<Tensor> r =
        launch(behaviour_logits.narrow(0, 1, T), learner_logits.narrow(0, 0, T), actions.narrow(0, 1, T),
               c10::nullopt, rewards.narrow(0, 1, T), learner_baseline.narrow(0, 0, T),
               learner_baseline.select(0, T), done.narrow(0, 1, T), o, want_grads);
    ctx->saved_data["dims"] = std::vector<int64_t>{T + 1, B, A};
    ctx->saved_data["need"] = std::vector<bool>{learner_logits.requires_grad(), learner_baseline.requires_grad()};
    ctx->saved_data["ran"] = false;
    return outputs_of(ctx, r);
  }

  static variable_list backward(AutogradContext* ctx, variable_list grad_outputs) {
    TORCH_CHECK(!ctx->saved_data["ran"].toBool(),
                "FusedLearnerLoss hands out its stashed gradients in place; backward can run only once per "
                "forward");
    ctx->saved_data["ran"] = true;
    const Tensor grads = ctx->saved_data["grads"].toTensor();
    const auto dims = ctx->saved_data["dims"].toIntVector();
    const auto need = ctx->saved_data["need"].toBoolList();
    variable_list out(14);
    if (grads.numel() == 0 || !grad_outputs[0].defined()) return out;
    scale_block(grads, grad_outputs[0]);
    if (need[0]) out[0] = d_logits_of(grads, dims[0], dims[1], dims[2]);
    if (need[1]) out[1] = d_values_of(grads, dims[0], dims[1], dims[2]);
    return out;
  }
};

// Pads the 4 outputs of a call without the aux fields to the 7 the Python side unpacks.
std::vector<c10::optional<Tensor>> as_optional(const variable_list& v) {
  std::vector<c10::optional<Tensor>> r;
  r.reserve(7);
  for (const Tensor& t : v) r.push_back(c10::optional<Tensor>(t));
  while (r.size() < 7) r.push_back(c10::nullopt);
  return r;
}

// Python: (total, losses, vs, pg_advantages, log_rhos | None, b_lp | None, t_lp | None)
std::vector<c10::optional<Tensor>> fused_loss(const Tensor& behaviour, const Tensor& target, const Tensor& actions,
                                              const c10::optional<Tensor>& discounts, const Tensor& rewards,
                                              const Tensor& values, const Tensor& bootstrap,
                                              const c10::optional<Tensor>& done, c10::optional<double> clip_rho,
                                              c10::optional<double> clip_pg_rho, double baseline_cost,
                                              double entropy_cost, double discounting, int64_t reward_clip_mode,
                                              bool want_aux, int64_t path) {
  return as_optional(FusedLossFn::apply(behaviour, target, actions, discounts, rewards, values, bootstrap, done,
                                        clip_rho, clip_pg_rho, baseline_cost, entropy_cost, discounting,
                                        reward_clip_mode, want_aux, path));
}

std::vector<c10::optional<Tensor>> fused_learner_loss(const Tensor& learner_logits, const Tensor& learner_baseline,
                                                      const Tensor& behaviour_logits, const Tensor& actions,
                                                      const Tensor& rewards, const Tensor& done,
                                                      c10::optional<double> clip_rho,
                                                      c10::optional<double> clip_pg_rho, double baseline_cost,
                                                      double entropy_cost, double discounting,
                                                      int64_t reward_clip_mode, bool want_aux, int64_t path) {
  return as_optional(FusedLearnerLossFn::apply(learner_logits, learner_baseline, behaviour_logits, actions, rewards,
                                               done, clip_rho, clip_pg_rho, baseline_cost, entropy_cost,
                                               discounting, reward_clip_mode, want_aux, path));
}

// The same two launches without an autograd node, for a learner that feeds the
// gradients to the heads' backward itself (torch.autograd.backward([logits,
// baseline], [d_logits, d_baseline])): the losses and both gradients are
// results of the one forward launch anyway, and the autograd engine's hop for
// this node costs more than the kernel at small shapes.
// Python: (total, losses, vs, pg_advantages, d_target_logits, d_values)
std::vector<Tensor> fused_loss_and_grads(const Tensor& behaviour, const Tensor& target, const Tensor& actions,
                                         const c10::optional<Tensor>& discounts, const Tensor& rewards,
                                         const Tensor& values, const Tensor& bootstrap,
                                         const c10::optional<Tensor>& done, c10::optional<double> clip_rho,
                                         c10::optional<double> clip_pg_rho, double baseline_cost,
                                         double entropy_cost, double discounting, int64_t reward_clip_mode,
                                         int64_t path) {
  const Options o = {clip_rho, clip_pg_rho, baseline_cost, entropy_cost, discounting, reward_clip_mode, path,
                     false, false};
  std::vector<Tensor> r = launch(behaviour, target, actions, discounts, rewards, values, bootstrap, done, o, true);
  const int64_t T = target.size(0), B = target.size(1), A = target.size(2);
  const Tensor grads = r.back();
  return {r[0], r[1], r[2], r[3], d_logits_of(grads, T, B, A), d_values_of(grads, T, B, A)};
}

// [T+1, ...] tensors as build_learner has them; gradients [T+1, ...] with a zero last step.
std::vector<Tensor> fused_learner_loss_and_grads(const Tensor& learner_logits, const Tensor& learner_baseline,
                                                 const Tensor& behaviour_logits, const Tensor& actions,
                                                 const Tensor& rewards, const Tensor& done,
                                                 c10::optional<double> clip_rho, c10::optional<double> clip_pg_rho,
                                                 double baseline_cost, double entropy_cost, double discounting,
                                                 int64_t reward_clip_mode, int64_t path) {
  TORCH_CHECK_VALUE(learner_logits.dim() == 3, "learner_policy_logits must have rank 3, got shape ",
                    learner_logits.sizes());
  const int64_t T = learner_logits.size(0) - 1, B = learner_logits.size(1), A = learner_logits.size(2);
  TORCH_CHECK_VALUE(T >= 1, "need at least 2 time steps (T+1), got ", T + 1);
  for (const Tensor* t : {&learner_baseline, &actions, &rewards, &done})
    TORCH_CHECK_VALUE(t->dim() == 2 && t->size(0) == T + 1 && t->size(1) == B,
                      "learner_baseline, actions, rewards and done must have shape [T+1, B]");
  TORCH_CHECK_VALUE(behaviour_logits.sizes() == learner_logits.sizes(),
                    "behaviour and learner logits differ in shape");
  const Options o = {clip_rho, clip_pg_rho, baseline_cost, entropy_cost, discounting, reward_clip_mode, path,
                     false, true};
  std::vector<Tensor> r =
      launch(behaviour_logits.narrow(0, 1, T), learner_logits.narrow(0, 0, T), actions.narrow(0, 1, T),
             c10::nullopt, rewards.narrow(0, 1, T), learner_baseline.narrow(0, 0, T),
             learner_baseline.select(0, T), done.narrow(0, 1, T), o, true);
  const Tensor grads = r.back();
  return {r[0], r[1], r[2], r[3], d_logits_of(grads, T + 1, B, A), d_values_of(grads, T + 1, B, A)};
}

// vtrace.from_importance_weights (vtrace.py:164-280) for the plain case: five
// contiguous float32 CUDA tensors, [T, B] and [B].  Everything else (trailing
// dimensions, other dtypes, host data) stays with the Python wrapper.
// Python: (vs, pg_advantages)
std::vector<Tensor> importance_weights(const Tensor& log_rhos, const Tensor& discounts, const Tensor& rewards,
                                       const Tensor& values, const Tensor& bootstrap,
                                       c10::optional<double> clip_rho, c10::optional<double> clip_pg_rho) {
  TORCH_CHECK(g_abi.importance_weights != nullptr, "vtrace_torch_ext: vtr_from_importance_weights_f32 not bound");
  const at::Device dev = values.device();
  want(values, "values", at::kFloat, 2, dev);
  want(log_rhos, "log_rhos", at::kFloat, 2, dev);
  want(discounts, "discounts", at::kFloat, 2, dev);
  want(rewards, "rewards", at::kFloat, 2, dev);
  want(bootstrap, "bootstrap_value", at::kFloat, 1, dev);
  const int64_t T = values.size(0), B = values.size(1);
  for (const Tensor* t : {&log_rhos, &discounts, &rewards})
    TORCH_CHECK_VALUE(t->size(0) == T && t->size(1) == B, "log_rhos, discounts and rewards must have shape [T, B] = [",
                      T, ", ", B, "]");
  TORCH_CHECK_VALUE(bootstrap.size(0) == B, "bootstrap_value must start with [B]");
  TORCH_CHECK_VALUE(T > 0 && B > 0 && T * B < (int64_t(1) << 31), "bad shape");
  c10::cuda::CUDAGuard guard(dev);
  Tensor out = at::empty({2, T, B}, values.options());
  const uint32_t flags = (clip_rho ? 0u : VTR_FLAG_NO_CLIP_RHO) | (clip_pg_rho ? 0u : VTR_FLAG_NO_CLIP_PG_RHO);
  cudaStream_t stream = at::cuda::getCurrentCUDAStream(dev.index()).stream();
  check(g_abi.importance_weights(log_rhos.data_ptr<float>(), discounts.data_ptr<float>(), rewards.data_ptr<float>(),
                                 values.data_ptr<float>(), bootstrap.data_ptr<float>(), static_cast<int32_t>(T),
                                 static_cast<int32_t>(B), 1, 0u, clip_rho ? static_cast<float>(*clip_rho) : 0.0f,
                                 clip_pg_rho ? static_cast<float>(*clip_pg_rho) : 0.0f, flags,
                                 out.data_ptr<float>(), out.data_ptr<float>() + T * B, stream),
        "vtr_from_importance_weights_f32");
  return {out.select(0, 0), out.select(0, 1)};
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("bind", &bind, "hand over the addresses of the C ABI entry points");
  m.def("fused_loss", &fused_loss);
  m.def("fused_learner_loss", &fused_learner_loss);
  m.def("bind_importance_weights", &bind_importance_weights);
  m.def("importance_weights", &importance_weights);
  m.def("fused_loss_and_grads", &fused_loss_and_grads);
  m.def("fused_learner_loss_and_grads", &fused_learner_loss_and_grads);
}
