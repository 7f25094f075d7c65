"""Pin the oracle against a REAL jax / flax install, if one ever appears -- TEST INFRASTRUCTURE.

    python oracle/check_against_jax.py [--dump DIR]

In this image `import jax` fails (no jax, jaxlib, flax or optax; no network), so the script prints that and exits 0: parity
of `oracle/reference_math.py` stays UNPINNED here.  Where jax imports, it evaluates the arithmetic of the reference scripts
with jax itself on forced host devices (XLA_FLAGS=--xla_force_host_platform_device_count=8, as case1a.py:3 sets it) and compares
with this repo's restatements, call site by call site:

  jax.random.PRNGKey / split / fold_in / bits / uniform / normal         vs oracle/jax_prng.py (integers bit-exact, floats 4 ulp)
  nn.initializers.lecun_normal()(key, (640, 512))                        vs jax_prng.lecun_normal            (needs flax)
  jax.lax.dot(A, B), A (4,16), B (16,4) f32, from one key                 vs reference_math.dot_f32           case1a.py:16-49
  jnp.einsum("ABC,ACD->ABD")                                              vs reference_math.einsum_abc_acd_abd  case4_gspmd_ff.py:30
  nn.Dense(512, use_bias=False, dtype=bf16) / Dense with bias            vs reference_math.dense_bf16        case6_attention.py:61-90
  the attention core written as in case6_attention.py:121-135             vs reference_math.attention_core
  jax.nn.gelu(approximate=True)                                           vs reference_math.gelu_tanh

With --dump every input and jax output is written as .npy under DIR (default oracle/_ref/jax_vectors), the material for
golden fixtures under tests/golden/ that would turn "parity unpinned" into a pinned oracle.  Exit status 1 on any mismatch.
"""
import argparse
import os
import sys

os.environ.setdefault("XLA_FLAGS", "--xla_force_host_platform_device_count=8")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _ulps(a, b):
    import numpy as np
    a, b = np.asarray(a, dtype=np.float32), np.asarray(b, dtype=np.float32)
    return np.abs(a.view(np.int32).astype(np.int64) - b.view(np.int32).astype(np.int64))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dump", default=None)
    args = ap.parse_args()
    try:
        import jax
        import jax.numpy as jnp
    except Exception as e:   # noqa: BLE001
        print(f"JAX not available in this image ({type(e).__name__}: {e}); the oracle stays unpinned against jax itself.")
        return 0
    import numpy as np
    import torch

    from oracle import jax_prng as jp
    from oracle import reference_math as rm

    dump_dir = args.dump
    if dump_dir:
        os.makedirs(dump_dir, exist_ok=True)

    def dump(name, arr):
        if dump_dir:
            np.save(os.path.join(dump_dir, name + ".npy"), np.asarray(arr))

    failures = []

    def check(name, ok, detail=""):
        print(f"{'ok  ' if ok else 'FAIL'} {name} {detail}")
        if not ok:
            failures.append(name)

    t = lambda a: torch.from_numpy(np.asarray(a, dtype=np.float32).copy())
    print(f"jax {jax.__version__}, {jax.device_count()} devices ({jax.devices()[0].platform})")

    # ---- PRNG
    key = jax.random.PRNGKey(0)
    kd = np.asarray(jax.random.key_data(key) if hasattr(jax.random, "key_data") else key, dtype=np.uint32)
    check("PRNGKey(0) key data", kd.tolist() == jp.PRNGKey(0).tolist(), str(kd.tolist()))
    sp = np.asarray(jax.random.split(key), dtype=np.uint32)
    check("split(PRNGKey(0))", sp.tolist() == jp.split(jp.PRNGKey(0)).tolist(), str(sp.tolist()))
    fi = np.asarray(jax.random.fold_in(key, 99), dtype=np.uint32)
    check("fold_in(PRNGKey(0), 99)", fi.tolist() == jp.fold_in(jp.PRNGKey(0), 99).tolist())
    for shape in ((1,), (3,), (4, 16), (7, 5, 3)):
        bits = np.asarray(jax.random.bits(key, shape, dtype=jnp.uint32))
        want = jp.random_bits(jp.PRNGKey(0), int(np.prod(shape))).reshape(shape)
        check(f"bits{shape}", np.array_equal(bits, want), "(differs if jax_threefry_partitionable is on in this jax)")
        nrm = np.asarray(jax.random.normal(key, shape))
        u = _ulps(nrm, jp.normal(jp.PRNGKey(0), shape)).max()
        check(f"normal{shape}", u <= 4, f"max {u} ulp")
        dump(f"normal_key0_{'x'.join(map(str, shape))}", nrm)
    uni = np.asarray(jax.random.uniform(key, (1000,), minval=-3.0, maxval=5.0))
    check("uniform(-3, 5)", _ulps(uni, jp.uniform(jp.PRNGKey(0), (1000,), -3.0, 5.0)).max() <= 1)
    try:
        import flax.linen as nn
        w = np.asarray(nn.initializers.lecun_normal()(jax.random.PRNGKey(1), (640, 512), jnp.float32))
        u = _ulps(w, jp.lecun_normal(jp.PRNGKey(1), (640, 512))).max()
        check("lecun_normal (640, 512)", u <= 8, f"max {u} ulp")
        dump("lecun_normal_key1_640x512", w)
    except ImportError:
        nn = None
        print("     flax not available: lecun_normal and nn.Dense comparisons skipped")

    # ---- case1a.py:16-49 -- A and B from ONE key, lax.dot in f32
    A = jax.random.normal(key, (4, 16))
    B = jax.random.normal(key, (16, 4))
    C = np.asarray(jax.lax.dot(A, B))
    want = rm.dot_f32(t(A), t(B)).numpy()
    check("lax.dot (4,16)@(16,4)", np.allclose(C, want, rtol=1e-5, atol=1e-6), f"max abs diff {np.abs(C - want).max():.2e}")
    for n, a in (("case1a_A", A), ("case1a_B", B), ("case1a_C", C)):
        dump(n, a)
    # ---- case4_gspmd_ff.py:26-33
    a3, b3 = jax.random.normal(key, (8, 4, 16)), jax.random.normal(jax.random.PRNGKey(1), (8, 16, 4))
    e = np.asarray(jnp.einsum("ABC,ACD->ABD", a3, b3))
    check("einsum ABC,ACD->ABD", np.allclose(e, rm.einsum_abc_acd_abd(t(a3), t(b3)).numpy(), rtol=1e-5, atol=1e-6))
    # ---- gelu
    xs = np.linspace(-6, 6, 4001, dtype=np.float32)
    g = np.asarray(jax.nn.gelu(jnp.asarray(xs), approximate=True))
    check("gelu(approximate=True)", np.allclose(g, rm.gelu_tanh(t(xs)).numpy(), rtol=1e-6, atol=1e-7))
    # ---- nn.Dense(dtype=bf16), case6_attention.py:61-90
    x = jax.random.normal(jax.random.PRNGKey(2), (2, 64, 640))
    if nn is not None:
        for use_bias in (False, True):
            layer = nn.Dense(512, use_bias=use_bias, dtype=jnp.bfloat16)
            params = layer.init(jax.random.PRNGKey(3), x)
            if use_bias:     # flax initialises the bias to zeros: give it values so that the bias path is exercised
                params = {"params": {"kernel": params["params"]["kernel"],
                                     "bias": 0.02 * jax.random.normal(jax.random.PRNGKey(4), (512,))}}
            y = np.asarray(layer.apply(params, x).astype(jnp.float32))
            wk = t(params["params"]["kernel"])
            wb = t(params["params"]["bias"]) if use_bias else None
            want = rm.dense_bf16(t(x), wk, wb).float().numpy()
            # identical roundings: at most one bf16 ulp where an f32 sum lands on a rounding boundary
            bad = np.abs(y - want) > 2.0 ** -7 * np.maximum(np.abs(want), 2.0 ** -10)
            check(f"nn.Dense(512, use_bias={use_bias}, dtype=bf16)", not bad.any(), f"{int(bad.sum())} of {bad.size} beyond one ulp")
            dump(f"dense_bias{int(use_bias)}_y", y)
    # ---- attention core as written at case6_attention.py:109-135
    q, k, v = (jax.random.normal(jax.random.PRNGKey(10 + i), (2, 128, 4, 64)).astype(jnp.bfloat16) for i in range(3))
    qf, kf = q.astype(jnp.float32), k.astype(jnp.float32)
    scores = jnp.einsum("b t n h, b f n h -> b n f t", kf, qf) * (64 ** -0.5)
    probs = jax.nn.softmax(scores, axis=-1).astype(jnp.bfloat16)
    o = np.asarray(jnp.einsum("b n f t, b t n h -> b f n h", probs, v).astype(jnp.float32))
    tb = lambda a: torch.from_numpy(np.asarray(a.astype(jnp.float32)).copy()).to(torch.bfloat16)
    want = rm.attention_core(tb(q), tb(k), tb(v)).float().numpy()
    bad = np.abs(o - want) > 2.0 ** -7 * np.maximum(np.abs(want), np.sqrt((want ** 2).mean()))
    check("attention core (case6_attention.py:121-135)", bad.mean() < 1e-3, f"{int(bad.sum())} of {bad.size} beyond one ulp")
    for n, a in (("attn_q", q.astype(jnp.float32)), ("attn_k", k.astype(jnp.float32)), ("attn_v", v.astype(jnp.float32)), ("attn_o", o)):
        dump(n, a)
    print("RESULT", "PASS" if not failures else "FAIL: " + ", ".join(failures))
    return 1 if failures else 0


if __name__ == "__main__":
    sys.exit(main())
