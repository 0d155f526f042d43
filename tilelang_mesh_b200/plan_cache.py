"""Measured, persisted tile plans.

The reference keeps what its autotuner found on disk and reuses it across processes
(tilelang/autotuner/tuner.py:281-287 `save_to_disk` / `load_from_disk`, tilelang/cache/kernel_cache.py:69-113:
a directory under ~/.tilelang/cache keyed by a hash of the program and its tunables).  Here every problem runs the
same kernel family and the tunables are the plan hints of the C ABI (`cfg_cta_group`, `cfg_block_n`, `cfg_split_k`,
`cfg_raster`, `cfg_stages`), so what is persisted is small: one JSON object per GPU model,

    { "<m>x<n>x<k>|<ab>-><c>|<layout>": {"cfg": {...}, "us": <median latency>, "auto_us": <latency of the cycle model's plan>,
                                         "candidates": <how many plans were timed>} }

`gemm.gemm()` consults it when it builds the descriptor of a problem for the first time and the caller gave no hints of
its own: a measured plan wins over the cycle model in `resolve()` (csrc/tlb200.cu).  Two layers: the read-only table
shipped in-tree (`plans/sm<count>.json`, produced by tools/tune_shapes.py on a B200) and the user's own
(`$TLB200_PLAN_CACHE`, default ~/.cache/tilelang_mesh_b200/plans_sm<count>.json), which `tune()` writes.
`TLB200_PLAN_CACHE=off` disables both.
"""
from __future__ import annotations

import json
import os
import tempfile
import threading
from typing import Dict, Iterable, Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
_lock = threading.Lock()
_tables: Dict[int, dict] = {}   # sm count -> merged table
_user_dirty: Dict[int, dict] = {}

_DT = {"torch.float16": "f16", "torch.bfloat16": "bf16", "torch.float8_e4m3fn": "e4m3", "torch.float8_e5m2": "e5m2",
       "torch.float32": "f32", "torch.int8": "i8", "torch.int32": "i32"}


def enabled() -> bool:
    return os.environ.get("TLB200_PLAN_CACHE", "").lower() not in ("off", "0", "none")


def key(m: int, n: int, k: int, ab_dtype, c_dtype, trans_a: bool, trans_b: bool) -> str:
    lay = ("T" if trans_a else "N") + ("T" if trans_b else "N")
    return f"{m}x{n}x{k}|{_DT.get(str(ab_dtype), str(ab_dtype))}->{_DT.get(str(c_dtype), str(c_dtype))}|{lay}"


def shipped_path(sm_count: int) -> str:
    return os.path.join(_HERE, "plans", f"sm{sm_count}.json")


def user_path(sm_count: int) -> str:
    p = os.environ.get("TLB200_PLAN_CACHE")
    if p and p.lower() not in ("off", "0", "none"):
        return p
    return os.path.join(os.path.expanduser("~"), ".cache", "tilelang_mesh_b200", f"plans_sm{sm_count}.json")


def _read(path: str) -> dict:
    try:
        with open(path) as f:
            d = json.load(f)
        return d if isinstance(d, dict) else {}
    except (OSError, ValueError):
        return {}


def table(sm_count: int) -> dict:
    with _lock:
        t = _tables.get(sm_count)
        if t is None:
            t = dict(_read(shipped_path(sm_count)))
            t.update(_read(user_path(sm_count)))  # the user's own measurements win
            _tables[sm_count] = t
        return t


def lookup(sm_count: int, m, n, k, ab_dtype, c_dtype, trans_a, trans_b) -> Optional[dict]:
    """The measured plan hints for this problem on this GPU model, or None."""
    if not enabled():
        return None
    e = table(sm_count).get(key(m, n, k, ab_dtype, c_dtype, trans_a, trans_b))
    return dict(e["cfg"]) if e and isinstance(e.get("cfg"), dict) else None


def store(sm_count: int, k: str, entry: dict, path: Optional[str] = None) -> str:
    """Record one measured plan in the user's table (atomic rewrite of the JSON file)."""
    path = path or user_path(sm_count)
    with _lock:
        os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
        cur = _read(path)
        cur[k] = entry
        fd, tmp = tempfile.mkstemp(dir=os.path.dirname(path) or ".", suffix=".tmp")
        with os.fdopen(fd, "w") as f:
            json.dump(cur, f, indent=0, sort_keys=True)
        os.replace(tmp, path)
        _tables.pop(sm_count, None)  # re-merge on next lookup
    return path


def reset():
    """Forget what was loaded (tests; after changing TLB200_PLAN_CACHE)."""
    with _lock:
        _tables.clear()


# ------------------------------------------------------------------------------------------------ measuring
def plan_space(m: int, n: int, k: int, es: int) -> Iterable[dict]:
    """Explicit plans worth timing for an (m, n, k) problem with `es`-byte operands: {1, 2, 2 x 64 rows, 4 CTAs per cluster} x
    tile width x K slices (x raster for many-tile problems).  `resolve()` rejects what is not legal for the shape."""
    yield {}
    widths = [w for w in (32, 64, 128, 256, 512) if w // 2 < n or w == 32]
    kb = -(-k * es // 128)
    for cg in (1, 2, 3, 4):
        if cg in (2, 4) and m <= 128 or cg == 3 and m <= 64:
            continue
        for bn in widths:
            if bn == 512 and (cg != 2 or es != 2):
                continue
            if cg == 4 and (bn > 256 or n <= bn):
                continue
            if cg == 3 and (bn % 64 or bn > 256):  # pair tiles of 128 rows
                continue
            tiles = -(-m // (128 if cg == 3 else 128 * min(cg, 2))) * -(-n // bn)
            splits = [1]
            if cg != 4 and tiles < 148 and kb >= 16:
                splits += [s for s in (2, 4, 8, 16) if kb // s >= 8 and tiles * s <= 2 * 148]
            for sk in splits:
                cfg = dict(cta_group=cg, block_n=bn, split_k=sk)
                yield cfg
                if tiles >= 4 * 148 and sk == 1 and bn >= 256:
                    for r in (4, 8, 16):
                        yield dict(cfg, raster=r)
    if -(-m // 256) * -(-n // 256) > 74:
        yield dict(split_k=-1)  # the tail split forced on the automatic tile


def tune(a, b, c, *, trans_a=False, trans_b=True, reps: int = 9, persist: bool = True, path: Optional[str] = None,
         candidates: Optional[Iterable[dict]] = None, verbose: bool = False) -> dict:
    """Time every plan of `plan_space` for exactly these tensors (CUDA events, L2 flushed before every launch, median
    of `reps`) and persist the winner.  Returns the table entry."""
    import torch

    from . import gemm as G
    m, ka = (a.shape[1], a.shape[0]) if trans_a else (a.shape[0], a.shape[1])
    n = b.shape[0] if trans_b else b.shape[1]
    dev = a.device
    cands = list(candidates) if candidates is not None else list(plan_space(m, n, ka, a.element_size()))

    # L2 is flushed by READING a buffer larger than it (a written flush buffer leaves the L2 full of dirty lines whose
    # write-back then competes with the first wave's loads), and the candidates are timed round-robin so that clock
    # and temperature drift hits them all alike.
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    G._plan_bypass = True  # time the plan named by cfg, not a cached one
    try:
        legal = []
        for cfg in cands:
            try:
                for _ in range(2):
                    G.gemm(a, b, c, trans_a=trans_a, trans_b=trans_b, **cfg)
                legal.append(cfg)
            except RuntimeError:
                continue  # not a legal plan for this shape
        if not legal:
            raise RuntimeError("plan_cache.tune: no plan could be timed")
        if legal and not legal[0] and len(legal) > 1:
            legal.append({})  # the cycle model's own plan is timed at both ends of every round: its spread is the noise
        # CUDA events resolve ~2 us: a launch shorter than ~50 us is timed as a group of back-to-back launches (the
        # first of them finds L2 flushed, the others do not - such a problem fits L2 and is normally called that way)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        G.gemm(a, b, c, trans_a=trans_a, trans_b=trans_b, **legal[0])
        e1.record()
        torch.cuda.synchronize(dev)
        first_us = max(e0.elapsed_time(e1) * 1e3, 2.0)
        inner = 1 if first_us >= 50.0 else min(64, int(200.0 / first_us) + 1)
        times = [[] for _ in legal]
        for _ in range(reps):
            for i, cfg in enumerate(legal):
                sink = flush.view(torch.int64).sum()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(inner):
                    G.gemm(a, b, c, trans_a=trans_a, trans_b=trans_b, **cfg)
                e1.record()
                torch.cuda.synchronize(dev)
                times[i].append(e0.elapsed_time(e1) * 1e3 / inner)
    finally:
        G._plan_bypass = False
    results = [(sorted(t)[len(t) // 2], cfg) for t, cfg in zip(times, legal)]
    autos = [us for us, cfg in results if not cfg]
    auto_us = min(autos) if autos else None
    noise = (max(autos) / min(autos) - 1.0) if len(autos) > 1 else 0.0
    best_us, best_cfg = min(results, key=lambda r: r[0])
    # A measured plan replaces the cycle model's only when it wins by 3 % AND by twice the spread between the two
    # timings of the identical automatic plan (on a 6 % spread an identical kernel once "won" by 6 %).
    if auto_us is not None and best_us > (1.0 - max(0.03, 2.0 * noise)) * auto_us:
        best_us, best_cfg = auto_us, {}
    entry = {"cfg": best_cfg, "us": round(best_us, 2), "auto_us": None if auto_us is None else round(auto_us, 2),
             "candidates": len(results), "auto_spread": round(noise, 4), "launches_per_sample": inner}
    if verbose:
        print(f"tune {m}x{n}x{ka}: best {best_cfg or 'automatic'} {best_us:.1f} us (automatic {auto_us and round(auto_us, 1)} us, "
              f"{len(results)} plans)")
    if persist and (best_cfg or path is None):  # a table file holds improvements over the cycle model only
        sm = torch.cuda.get_device_properties(dev).multi_processor_count
        store(sm, key(m, n, ka, a.dtype, c.dtype, trans_a, trans_b), entry, path)
    return entry
