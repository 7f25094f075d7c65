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
