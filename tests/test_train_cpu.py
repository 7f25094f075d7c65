"""train_step of the reference's single-Dense module (case5_attention_dense.py:130-139)
on a GPU-less box: the oracle against closed forms, and the sharded plan of the host layer (virtual 2x2 mesh, per-shard
compute routed to the oracle-backed checker) against the unsharded oracle."""
import pytest
import torch

from oracle import reference_math as rm


def test_oracle_first_adam_step_is_signed_learning_rate():
    """With zero moments the first optax.adam update is -lr * g / (|g| + eps): the bias corrections cancel b1, b2."""
    x, w = rm.synthetic((2, 16, 32), 1), rm.synthetic((32, 24), 2, "weight")
    w2, mu, nu, g = rm.dense_train_step(x, w, torch.zeros_like(w), torch.zeros_like(w), 1)
    torch.testing.assert_close(w2 - w, -1e-3 * g / (g.abs() + 1e-8), rtol=0, atol=3e-8)   # f32 rounding of w - lr*u
    torch.testing.assert_close(mu, 0.1 * g, rtol=1e-6, atol=0)
    torch.testing.assert_close(nu, 0.001 * g * g, rtol=1e-4, atol=0)      # (1 - 0.999) evaluated in f32, as optax does


def test_oracle_dense_gradient_is_bf16_column_sum():
    """d sum(x.W) / dW[k, n] = sum over rows of bf16(x)[., k] for every n, rounded once to bf16."""
    x, w = rm.synthetic((3, 8, 16), 3), rm.synthetic((16, 8), 4, "weight")
    g, _ = rm.dense_sum_loss_grads(x, w)
    col = x.to(torch.bfloat16).double().reshape(-1, 16).sum(0)
    want = col.float().to(torch.bfloat16).float()
    assert torch.equal(g, want[:, None].expand(16, 8))
    # a finite-difference probe of the f64 loss agrees
    eps = 1e-3
    w64 = w.double()
    base = (x.to(torch.bfloat16).double().reshape(-1, 16) @ w64).sum()
    w64[5, 2] += eps
    fd = ((x.to(torch.bfloat16).double().reshape(-1, 16) @ w64).sum() - base) / eps
    assert abs(fd.item() - col[5].item()) < 1e-6 * max(1.0, abs(col[5].item()))


@pytest.mark.parametrize("wspec", [("embed", "kv"), ("embed", "hidden"), (None, "hidden")])
def test_sharded_dense_train_step_matches_oracle(ljs, cpu_checker, wspec):
    """case5's layout: x P('data','model'), kernel by the script's rules; two train steps."""
    rt = ljs.init_virtual(4, "cpu")
    mesh = ljs.Mesh(ljs.create_device_mesh((2, 2)), ("data", "model"))
    P = ljs.PartitionSpec
    rules = (("batch", "data"), ("embed", "data"), ("hidden", "model"))      # case5_attention_dense.py:109-112
    x_host, w_host = rm.synthetic((4, 32, 64), 1), rm.synthetic((64, 48), 2, "weight")
    x = ljs.device_put(x_host, ljs.NamedSharding(mesh, P("data", "model")))
    w = ljs.device_put(w_host.clone(), ljs.nn.logical_to_mesh_sharding(wspec, mesh, rules))
    state = ljs.train.TrainState.create(apply_fn=None, params={"kernel": w}, tx=ljs.train.adam(learning_rate=0.001))
    rt.collective_log.clear()
    g = ljs.train.dense_sum_loss_grad(x, w)
    assert [n for n, _, _ in rt.collective_log][:1] == ["all_reduce"]        # the data-parallel gradient reduction
    assert g.sharding.same_layout(w.sharding, w.shape)
    g_want, _ = rm.dense_sum_loss_grads(x_host, w_host)
    torch.testing.assert_close(g.global_array(), g_want, rtol=2 ** -7, atol=0)
    k, mu, nu = w_host.clone(), torch.zeros_like(w_host), torch.zeros_like(w_host)
    for step in (1, 2):
        state = ljs.train.dense_train_step(state, x)
        k, mu, nu, _ = rm.dense_train_step(x_host, k, mu, nu, step)
        assert state.step == step
        torch.testing.assert_close(state.params["kernel"].global_array(), k, rtol=0, atol=2e-5)
        torch.testing.assert_close(state.mu["kernel"].global_array(), mu, rtol=2 ** -6, atol=1e-6)
    # the optimizer state keeps the parameters' sharding (what logical_to_mesh_sharding gives the reference's state)
    assert state.mu["kernel"].sharding.same_layout(w.sharding, w.shape)


def test_train_state_rejects_non_f32_params(ljs, cpu_checker):
    ljs.init_virtual(2, "cpu")
    mesh = ljs.Mesh(ljs.create_device_mesh((2,)), ("data",))
    w = ljs.device_put(torch.zeros(8, 8, dtype=torch.bfloat16), ljs.NamedSharding(mesh, ljs.PartitionSpec()))
    with pytest.raises(TypeError):
        ljs.train.TrainState.create(apply_fn=None, params={"kernel": w}, tx=ljs.train.adam())


@pytest.mark.parametrize("causal", [False, True])
def test_attention_module_train_step_matches_oracle(ljs, cpu_checker, causal):
    """train_step of the attention module (case6_attention.py:206-220) in the head-sharded layout on a virtual 2x2 mesh:
    gradients in the parameters' shardings against autodiff of the unsharded f64 formula, then one Adam step."""
    from tests.test_plans_cpu import _attention_setup
    mesh, model, host, rules, x_host, x_sharding = _attention_setup(ljs, "heads", causal=causal)
    rt = ljs.get_runtime()
    params = model.shard_params(host, mesh, rules)
    x = ljs.device_put(x_host, x_sharding)
    want = rm.mha_sum_loss_grads(x_host, host, model.heads, model.dim_head, causal)
    with mesh, ljs.nn.axis_rules(rules):
        rt.collective_log.clear()
        grads = ljs.train.attention_sum_loss_grads(model, params, x)
        names = [n for n, _, _ in rt.collective_log]
        assert names.count("all_reduce") == 4                                 # dWq, dWk, dWv, dWo over the data axis
        for name, g in grads.items():
            target = params["to_out_0"][0 if name.endswith("kernel") else 1] if name.startswith("to_out_0") else params[name]
            assert g.sharding.same_layout(target.sharding, target.shape), name
            scale = want[name].abs().max().item()
            torch.testing.assert_close(g.global_array(), want[name], rtol=3e-2, atol=3e-2 * scale, msg=lambda m, n=name: n + ": " + m)
        state = ljs.train.attention_train_state(model, params, ljs.train.adam(learning_rate=0.001))
        state = ljs.train.attention_train_step(model, state, x)
    assert state.step == 1
    for name in ("to_q", "to_out_0.kernel"):
        host_p = host[name] if name == "to_q" else host["to_out_0"][0]
        k2, _, _ = rm.adam_update(host_p, want[name], torch.zeros_like(host_p), torch.zeros_like(host_p), 1)
        # the first Adam step is -lr * sign(g): compare where the gradient is well away from zero (its sign is then settled)
        mask = want[name].abs() > 0.1 * want[name].abs().max()
        torch.testing.assert_close(state.params[name].global_array()[mask], k2[mask], rtol=0, atol=2e-5)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("mesh_shape", [(2, 2), (1, 4)])
def test_attention_module_train_step_literal_annotations(ljs, cpu_checker, causal, mesh_shape):
    """The same train_step under the reference's LITERAL annotation set (activations split along S over 'model', q/k/v
    kernels split on their rows, to_out_0 on its columns -- case6_attention.py:56-59,85-88,105-116,183-187): gradients in the
    parameters' shardings against autodiff of the unsharded formula; the backward's collectives are the transposes of the
    forward's resharding points (reduce-scatter along S for dO, dK, dV; all-reduce of the weight gradients)."""
    from tests.test_plans_cpu import _attention_setup
    mesh, model, host, rules, x_host, x_sharding = _attention_setup(ljs, "literal", mesh_shape=mesh_shape, causal=causal)
    rt = ljs.get_runtime()
    params = model.shard_params(host, mesh, rules)
    assert tuple(params["to_q"].device_buffers[0].shape) == (64 // mesh_shape[1], 64)
    x = ljs.device_put(x_host, x_sharding)
    want = rm.mha_sum_loss_grads(x_host, host, model.heads, model.dim_head, causal)
    with mesh, ljs.nn.axis_rules(rules):
        rt.collective_log.clear()
        grads = ljs.train.attention_sum_loss_grads(model, params, x)
        names = [n for n, _, _ in rt.collective_log]
        assert names.count("reduce_scatter") == 3                              # dO, dK, dV along S over 'model'
        for name, g in grads.items():
            target = params["to_out_0"][0 if name.endswith("kernel") else 1] if name.startswith("to_out_0") else params[name]
            assert g.sharding.same_layout(target.sharding, target.shape), name
            scale = want[name].abs().max().item()
            torch.testing.assert_close(g.global_array(), want[name], rtol=3e-2, atol=3e-2 * scale, msg=lambda m, n=name: n + ": " + m)
        state = ljs.train.attention_train_state(model, params, ljs.train.adam(learning_rate=0.001))
        state = ljs.train.attention_train_step(model, state, x)
        y = model(ljs.train.attention_params_of(state), x)                     # the updated parameters run forward
    assert state.step == 1 and y.shape == x.shape
    moved = (state.params["to_q"].global_array() - host["to_q"]).abs()
    assert 0.5e-3 < moved.mean().item() <= 1.01e-3                             # first Adam step: |update| = lr almost everywhere
