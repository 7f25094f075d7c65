"""``jax.debug.visualize_array_sharding`` as the case scripts call it after every placement (case1a.py:26,32,51;
case5_attention_dense.py:121): a text picture of which devices hold which block of a 1-D or 2-D array.  Diagnostics only
(SURVEY.md section 8(f) row 4); plain text instead of the reference's coloured table, same information: one cell per
distinct block, labelled with the ids of the devices that hold it."""
from __future__ import annotations

from typing import Dict, List, Tuple

from .array import ShardedArray


def _blocks(arr: ShardedArray):
    """{(row slice, col slice) as (start, stop) tuples: sorted device ids} over ALL devices of the sharding."""
    index = arr.sharding.devices_indices_map(arr.shape)
    out: Dict[Tuple[Tuple[int, int], ...], List[int]] = {}
    for dev, idx in index.items():
        key = tuple((sl.start or 0, sl.stop if sl.stop is not None else dim) for sl, dim in zip(idx, arr.shape))
        out.setdefault(key, []).append(dev)
    return {k: sorted(v) for k, v in out.items()}


def sharding_text(arr: ShardedArray, platform: str = "GPU", max_width: int = 100) -> str:
    if arr.ndim not in (1, 2):
        raise ValueError(f"can only visualize 1-D or 2-D arrays, got rank {arr.ndim}")
    blocks = _blocks(arr)
    if arr.ndim == 1:
        blocks = {((0, 1), k[0]): v for k, v in blocks.items()}
    rows = sorted({k[0] for k in blocks})
    cols = sorted({k[1] for k in blocks})
    label = {k: f"{platform} " + ",".join(str(d) for d in v) for k, v in blocks.items()}
    width = max(len(t) for t in label.values()) + 2
    width = max(width, min(max_width // max(len(cols), 1), width))
    line = "+" + "+".join("-" * width for _ in cols) + "+"
    text = [line]
    for r in rows:
        text.append("|" + "|".join(label.get((r, c), "").center(width) for c in cols) + "|")
        text.append(line)
    return "\n".join(text)


def visualize_array_sharding(arr: ShardedArray, platform: str = "GPU") -> None:
    print(sharding_text(arr, platform))
