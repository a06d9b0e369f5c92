"""stats.py on the device: same numbers as on the host (the reductions are torch ops on whichever device holds the data)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_statistics_and_recorder_on_device():
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from sevennet_finetuning_b200 import stats
    from test_stats_cpu import NSP, _METRICS, _fake_output, _graph, _naive
    graphs = [_graph(5 + s % 4, 20 + s, s) for s in range(40)]
    a, b = stats.DatasetStatistics(graphs, NSP), stats.DatasetStatistics(graphs, NSP, device='cuda:0')
    assert abs(a.get_avg_num_neigh() - b.get_avg_num_neigh()) < 1e-12
    assert abs(a.get_force_rms() - b.get_force_rms()) < 1e-7
    assert torch.allclose(a.get_species_wise_force_rms(), b.get_species_wise_force_rms().cpu(), atol=1e-6)
    assert np.abs(a.get_species_ref_energy_by_linear_comb() - b.get_species_ref_energy_by_linear_comb()).max() < 1e-9
    rec = stats.ErrorRecorder(_METRICS, device='cuda:0')
    outs = [_fake_output(s) for s in range(4)]
    for o in outs:
        rec.update({k: v.cuda() for k, v in o.items()})
    got, want = rec.epoch_forward(), _naive(outs)
    for k in want:
        assert abs(got[k] - want[k]) <= 1e-6 * max(1.0, abs(want[k])), k
