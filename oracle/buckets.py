"""Aspect-ratio bucket assignment (oracle; test infrastructure only)."""


def find_closest_ratio_trainer(aspect_ratios: dict, ratio: float) -> str:
    """common/trainer.py:109-120 — first key in dict order minimising |float(key)-ratio|,
    strict ``>`` so ties keep the earlier key; initial distance 100, default '0.6'."""
    min_distance = 100
    target_ratio = 0.6
    for r in aspect_ratios.keys():
        distance = abs(float(r) - ratio)
        if min_distance > distance:
            target_ratio = r
            min_distance = distance
    return str(target_ratio)


def find_closest_ratio_sampler(aspect_ratios: dict, ratio: float) -> float:
    """common/bucket_sampler.py:314-322 — same argmin, initial error 1000, default -1,
    returns float(key)."""
    error = 1000
    res = -1
    for r in aspect_ratios.keys():
        new_error = abs(float(r) - ratio)
        if new_error < error:
            error = new_error
            res = float(r)
    return res


def find_closest_key(bucket_keys, ratio: float) -> float:
    """common/bucket_sampler.py:184-195 — argmin over the live bucket keys (floats)."""
    if ratio == -1.0:
        return -1.0
    error = 1000
    res = -1.0
    for r in bucket_keys:
        new_error = abs(ratio - r)
        if new_error < error:
            error = new_error
            res = r
    return res


def bucket_ready(counts_per_rank, batch_size: int) -> bool:
    """common/bucket_sampler.py:229-237 — a bucket is consumed only when every rank holds
    at least ``batch_size`` samples of that key."""
    return min(counts_per_rank) >= batch_size
