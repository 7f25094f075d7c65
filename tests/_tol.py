"""Error bounds for the bf16 parity tests: tolerances that scale with the size of the expected output.

BASELINE.json's north_star states "bf16 rtol 2e-2 with an fp32-accumulate check".  A flat atol next to that rtol is only
harmless while the outputs are O(1); attention outputs over Sk >= 1024 random keys are ~0.03, so the absolute term here is
tied to the row's own magnitude instead:

  tier T1   |got - T1|  <=  rtol * |T1| + c * 2^-8 * rms_row(T1)
            T1 = the oracle with bf16 roundings where the reference rounds; rms_row = root mean square over the last
            (feature / head-dim) axis of the same row, floored at 5 % of the tensor's overall rms so that rows that are
            exactly zero (causal row 0 of dQ) do not demand bit equality of two different summation orders;
  tier F64  |got - f64| <=  |T1 - f64| + 2 bf16 ulp,   ulp = 2^-7 * max(|f64|, rms_row)
            the analogue of the GEMM one-ulp check: the kernel may not be further from exact math than the rounding
            oracle is, plus two units in the last place of a bf16 number of the row's magnitude.

Gradients additionally pass `terms` (oracle attention_core_grad_term_scales): the row scale becomes max(rms_row(want),
rms_row(terms)).  The two agree except where a sum cancels: row f of a causal dQ is scale * sum_t P (dP - sum P dP) K
over f + 1 keys, and for the first few rows the bracket is a difference of O(10) numbers that leaves O(0.1); any bf16
pipeline -- this kernel (bf16 P, dS, O) as much as the reference's own autodiff (bf16 dP) -- then carries an absolute
error of 2^-9 x the size of the terms, which an f64 oracle does not have.
"""
import torch


def row_rms(t: torch.Tensor) -> torch.Tensor:
    t = t.double()
    rms = t.pow(2).mean(dim=-1, keepdim=True).sqrt()
    floor = 0.05 * t.pow(2).mean().sqrt()
    return torch.maximum(rms, floor.expand_as(rms)).clamp_min(1e-30)


def _row_scale(want: torch.Tensor, terms) -> torch.Tensor:
    """rms_row(want), or, where a `terms` tensor is given (the root-sum-square of the terms each element is a sum of --
    oracle attention_core_grad_term_scales), the larger of the two row magnitudes: rows whose sum cancels (the first
    rows of a causally masked dQ) are measured by what was added up, not by what is left."""
    scale = row_rms(want)
    if terms is not None:
        scale = torch.maximum(scale, row_rms(terms))
    return scale


def scaled_error(got: torch.Tensor, want: torch.Tensor, rtol: float = 2e-2, terms=None) -> float:
    """The c that tier T1 would need: max over elements of (|got - want| - rtol |want|) / (2^-8 row scale)."""
    got, want = got.double(), want.double()
    excess = ((got - want).abs() - rtol * want.abs()).clamp_min(0.0)
    return float((excess / (2.0 ** -8 * _row_scale(want, terms))).max())


def assert_close_scaled(got, want, *, rtol: float = 2e-2, c: float = 4.0, what: str = "", terms=None):
    """Tier T1."""
    assert torch.isfinite(got.double()).all(), f"{what}: non-finite output"
    need = scaled_error(got, want, rtol, terms)
    assert need <= c, (f"{what}: |got - want| exceeds {rtol} |want| + {c} * 2^-8 * rms_row by a factor "
                       f"{need / c:.2f} (needed c = {need:.2f}; max abs err {(got.double() - want.double()).abs().max():.3e})")


def f64_tier_ulps(got, t1, f64, terms=None) -> float:
    got, t1, f64 = got.double(), t1.double(), f64.double()
    ulp = 2.0 ** -7 * torch.maximum(f64.abs(), _row_scale(f64, terms))
    return float((((got - f64).abs() - (t1 - f64).abs()) / ulp).max())


def assert_no_further_from_f64(got, t1, f64, *, ulps: float = 2.0, what: str = "", terms=None):
    """Tier F64."""
    need = f64_tier_ulps(got, t1, f64, terms)
    assert need <= ulps, f"{what}: kernel is {need:.2f} bf16 ulp further from the f64 result than the rounding oracle (allowed {ulps})"
