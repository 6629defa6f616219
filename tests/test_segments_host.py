"""Host logic of diffsinger_b200.segments: how batch-1 segments are grouped into ragged batches (no GPU)."""
from diffsinger_b200.segments import _ragged_groups

CONFIG1 = [504, 522, 539, 720, 1197, 550, 550, 515, 550, 539, 539, 540, 1917, 549, 550, 515, 687, 549, 549, 549, 460, 286, 578, 545, 515,
           566, 572, 542, 516, 579]  # frames of the 30 segments of samples/00


def _check(groups, lengths, cap, min_fill):
    seen = sorted(i for g in groups for i in g)
    assert seen == list(range(len(lengths))), "every segment exactly once"
    for g in groups:
        t_max = max(lengths[i] for i in g)
        assert lengths[g[0]] == t_max, "longest first"
        assert len(g) == 1 or len(g) * -(-t_max // 128) <= cap, (g, cap)
        assert all(lengths[i] >= min_fill * t_max for i in g)


def test_whole_network_kernel_grouping_respects_the_cluster_capacity():
    groups = _ragged_groups(CONFIG1, 33, 0.0, 33, 0)  # every launch costs the same: fewest groups
    _check(groups, CONFIG1, 33, 0.0)
    assert len(groups) == 6  # 30 segments in six launches per evaluation
    assert sum(len(g) * -(-CONFIG1[g[0]] // 128) for g in groups) <= 6 * 33


def test_per_layer_grouping_trades_padding_against_call_overhead():
    groups = _ragged_groups(CONFIG1, 512, 0.25, 0, 62)  # a sampler call costs as much as 62 tiles of work
    _check(groups, CONFIG1, 512, 0.25)
    assert len(groups) == 2
    padded = sum(len(g) * CONFIG1[g[0]] for g in groups)
    assert padded < 1.5 * sum(CONFIG1), padded
    cheap = _ragged_groups(CONFIG1, 512, 0.25, 0, 8)     # cheap calls: more groups, less padding
    _check(cheap, CONFIG1, 512, 0.25)
    assert len(cheap) > len(groups)
    assert sum(len(g) * CONFIG1[g[0]] for g in cheap) < padded


def test_degenerate_inputs():
    assert _ragged_groups([], 33) == []
    assert _ragged_groups([100], 33) == [[0]]
    groups = _ragged_groups([5000, 100], 33)  # a segment bigger than the capacity still gets its own batch
    assert groups == [[0], [1]]
