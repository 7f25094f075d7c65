"""The reference's only machine-checkable pins: its ten shard-shape asserts, plus the
device -> block tables of SURVEY.md §3.1 and the case6 sharding table of §3.2."""
import numpy as np
import pytest
import torch


@pytest.fixture
def mesh8(ljs):
    ljs.init_virtual(8, "cpu")
    return ljs.PositionalSharding(ljs.create_device_mesh((2, 4)))


def _ab():
    g = torch.Generator().manual_seed(0)
    return torch.randn(4, 16, generator=g), torch.randn(16, 4, generator=g)


def test_case1a_shard_shapes(ljs, mesh8):
    A, B = _ab()
    a = ljs.device_put(A, mesh8.replicate(axis=0, keepdims=True))            # case1a.py:24
    b = ljs.device_put(B, mesh8.reshape(4, 2).replicate(axis=1, keepdims=True))   # case1a.py:30
    assert tuple(a.device_buffers[0].shape) == (4, 4)                        # case1a.py:36
    assert tuple(b.device_buffers[0].shape) == (4, 4)                        # case1a.py:43
    assert torch.equal(a.device_buffers[0], a.device_buffers[4])             # case1a.py:40 prints True
    assert torch.equal(b.device_buffers[0], b.device_buffers[1])             # case1a.py:46 (B_0 vs B_1)
    # device d holds A's K block d % 4 and B's K block d // 2
    for d in range(8):
        assert a.index_of(d)[1] == slice(4 * (d % 4), 4 * (d % 4) + 4)
        assert b.index_of(d)[0] == slice(4 * (d // 2), 4 * (d // 2) + 4)


def test_case1b_shard_shapes(ljs, mesh8):
    A, B = _ab()
    a = ljs.device_put(A, mesh8.replicate(axis=0, keepdims=True))            # case1b.py:24
    b = ljs.device_put(B, mesh8.replicate(axis=1, keepdims=True))            # case1b.py:30
    assert tuple(a.device_buffers[0].shape) == (4, 4)                        # case1b.py:36
    assert tuple(b.device_buffers[0].shape) == (8, 4)                        # case1b.py:42
    assert torch.equal(b.device_buffers[0], b.device_buffers[1])


def test_case2_shard_shapes(ljs, mesh8, cpu_checker):
    A, B = _ab()
    a = ljs.device_put(A, mesh8)                                             # case2.py:23
    b = ljs.device_put(B, mesh8.replicate(axis=1, keepdims=True))            # case2.py:29
    assert tuple(a.device_buffers[0].shape) == (2, 4)
    assert not torch.equal(a.device_buffers[0], a.device_buffers[4])
    c = ljs.lax.dot(a, b)                                                    # case2.py:44
    assert tuple(c.device_buffers[0].shape) == (2, 4)                        # case2.py:52
    assert torch.equal(c.device_buffers[0], c.device_buffers[1])             # case2.py:58
    assert not torch.equal(c.device_buffers[0], c.device_buffers[4])


def test_case3_shard_shapes(ljs, mesh8, cpu_checker):
    A, B = _ab()
    a = ljs.device_put(A, mesh8)                                             # case3_fully_sharded.py:23
    b = ljs.device_put(B, mesh8)                                             # case3_fully_sharded.py:29
    assert tuple(a.device_buffers[0].shape) == (2, 4)                        # :35
    assert tuple(b.device_buffers[0].shape) == (8, 1)                        # :41
    c = ljs.lax.dot(a, b)                                                    # :46
    assert tuple(c.device_buffers[0].shape) == (2, 1)                        # :52


def test_case4_shard_shapes(ljs, mesh8, cpu_checker):
    g = torch.Generator().manual_seed(0)
    arr_a, arr_b = torch.randn(8, 4, 16, generator=g), torch.randn(8, 16, 4, generator=g)
    c3 = ljs.lax.einsum("ABC,ACD->ABD", ljs.device_put(arr_a), ljs.device_put(arr_b))   # case4_gspmd_ff.py:30
    assert c3.shape == (8, 4, 4)                                             # case4_gspmd_ff.py:32
    A, B = _ab()
    a = ljs.device_put(A, mesh8.replicate(axis=1, keepdims=True))            # case4_gspmd_ff.py:46
    b = ljs.device_put(B, mesh8.replicate(axis=0, keepdims=True))            # case4_gspmd_ff.py:49
    assert tuple(a.device_buffers[0].shape) == (2, 16)
    assert tuple(b.device_buffers[0].shape) == (16, 1)
    c = ljs.lax.dot(a, b)                                                    # case4_gspmd_ff.py:52
    assert tuple(c.addressable_shards[0].data.shape) == (2, 1)               # case4_gspmd_ff.py:56-57
    assert ljs.lax.plan_dot(a, b)["strategy"] == "local"                     # zero-communication layout


def test_case6_sharding_table(ljs):
    """per-device shard shapes under the literal rules on a 2x2 mesh."""
    ljs.init_virtual(4, "cpu")
    mesh = ljs.Mesh(ljs.create_device_mesh((2, 2)), ("data", "model"))       # case6_attention.py:155-156
    P = ljs.PartitionSpec
    rules = (("batch", "data"), ("embed", "model"), ("hidden", "model"))     # case6_attention.py:183-187
    x = ljs.device_put(torch.zeros(8, 256, 640), ljs.NamedSharding(mesh, P("data", "model")))   # :161-162
    assert tuple(x.device_buffers[0].shape) == (4, 128, 640)
    to_q = ljs.nn.logical_to_mesh_sharding(("embed", "heads"), mesh, rules)  # :56-59
    assert to_q.spec == P("model", None)
    assert to_q.shard_shape((640, 512)) == (320, 512)
    to_out = ljs.nn.logical_to_mesh_sharding(("heads", "embed"), mesh, rules)   # :85-88
    assert to_out.spec == P(None, "model")
    assert to_out.shard_shape((512, 640)) == (512, 320)
    with ljs.nn.axis_rules(rules):
        assert ljs.nn.logical_to_mesh_axes(("batch", "embed", None)) == P("data", "model", None)      # :105-107
        assert ljs.nn.logical_to_mesh_axes(("batch", "embed", None, None)) == P("data", "model", None, None)
        assert ljs.nn.logical_to_mesh_axes(("batch", "kv", "heads")) == P("data", None, None)         # :137
        assert ljs.nn.logical_to_mesh_axes(("batch", "embed")) == P("data", "model")                  # :141
    # case5 rules: to_q kernel ('embed','kv') -> P('data', None), shard (320, 512)
    rules5 = (("batch", "data"), ("embed", "data"), ("hidden", "model"))     # case5_attention_dense.py:109-112
    s5 = ljs.nn.logical_to_mesh_sharding(("embed", "kv"), mesh, rules5)
    assert s5.spec == P("data", None) and s5.shard_shape((640, 512)) == (320, 512)
    # a mesh axis is used at most once per array
    with ljs.nn.axis_rules(rules):
        assert ljs.nn.logical_to_mesh_axes(("embed", "hidden")) == P("model", None)


def test_named_sharding_short_spec_and_replicated(ljs):
    ljs.init_virtual(8, "cpu")
    mesh = ljs.Mesh(ljs.create_device_mesh((2, 4)), ("data", "model"))
    s = ljs.NamedSharding(mesh, ljs.PartitionSpec("data"))
    assert s.shard_shape((8, 6, 10)) == (4, 6, 10)
    r = ljs.NamedSharding(mesh, None)
    assert r.is_fully_replicated((3, 5))
    both = ljs.NamedSharding(mesh, ljs.PartitionSpec(("data", "model"), None))
    assert both.shard_shape((16, 3)) == (2, 3)
    idx = both.devices_indices_map((16, 3))
    assert idx[5][0] == slice(10, 12)            # device (1,1): part 1*4+1 = 5
    assert mesh.axis_groups(("model",)) == [[0, 1, 2, 3], [4, 5, 6, 7]]
    assert mesh.axis_groups(("data",)) == [[0, 4], [1, 5], [2, 6], [3, 7]]
    with pytest.raises(ValueError):
        ljs.NamedSharding(mesh, ljs.PartitionSpec("data", "data"))
    with pytest.raises(ValueError):
        s.shard_shape((7, 2))


def test_visualize_array_sharding_text(ljs):
    """jax.debug.visualize_array_sharding (case1a.py:26,32,51): one cell per distinct block, labelled with its devices."""
    ljs.init_virtual(8, "cpu")
    s = ljs.PositionalSharding(ljs.create_device_mesh((2, 4)))
    import torch
    a = ljs.device_put(torch.zeros(4, 16), s.replicate(axis=0, keepdims=True))        # case1a.py:24: columns over 4, replicated over 2
    text = ljs.debug.sharding_text(a, platform="CPU")
    cells = [c.strip() for c in text.splitlines()[1].strip("|").split("|")]
    assert cells == ["CPU 0,4", "CPU 1,5", "CPU 2,6", "CPU 3,7"]
    b = ljs.device_put(torch.zeros(16, 4), s.reshape(4, 2).replicate(axis=1, keepdims=True))   # case1a.py:30
    rows = [ln for ln in ljs.debug.sharding_text(b, platform="CPU").splitlines() if ln.startswith("|")]
    assert [r.strip("|").strip() for r in rows] == ["CPU 0,1", "CPU 2,3", "CPU 4,5", "CPU 6,7"]
    mesh = ljs.Mesh(ljs.create_device_mesh((2, 4)), ("data", "model"))
    v = ljs.device_put(torch.zeros(8), ljs.NamedSharding(mesh, ljs.PartitionSpec("data")))
    assert "CPU 0,1,2,3" in ljs.debug.sharding_text(v, platform="CPU")
    import pytest
    with pytest.raises(ValueError):
        ljs.debug.sharding_text(ljs.device_put(torch.zeros(2, 2, 2), ljs.NamedSharding(mesh, ljs.PartitionSpec())))
