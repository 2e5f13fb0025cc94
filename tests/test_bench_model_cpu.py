"""bench.py's byte model against the figures SURVEY.md §8(d) derives for BASELINE.json's configs: the roofline fractions
the bench line reports are only as good as these denominators, so they are pinned here (no GPU needed)."""
import importlib.util
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bench)

from distributed_inference_server_b200 import LLAMA3_8B, LLAMA3_70B, TINYLLAMA  # noqa: E402


def step_bytes(shape, tp, batch, ctx):
    w, kv = bench.model_bytes(shape, tp)
    return w + batch * ctx * kv


def test_streamed_parameters_and_kv_bytes_per_token():
    w8, kv8 = bench.model_bytes(LLAMA3_8B, 1)
    assert w8 / 2 == pytest.approx(7.5049e9, rel=1e-4)            # P_stream, SURVEY.md §8(d)
    assert kv8 == 128 * 1024                                      # 128 KiB of bf16 K/V per token of context
    w70, kv70 = bench.model_bytes(LLAMA3_70B, 1)
    assert w70 / 2 == pytest.approx(69.503e9, rel=1e-4) and kv70 == 320 * 1024
    wt, kvt = bench.model_bytes(TINYLLAMA, 1)
    assert wt / 2 == pytest.approx(1.0345e9, rel=1e-3) and kvt == 22 * 1024


@pytest.mark.parametrize("tp,batch,want_gb", [(1, 1, 15.08), (1, 32, 17.16), (1, 64, 19.31), (2, 64, 9.65), (4, 64, 4.83), (8, 64, 2.41)])
def test_bytes_per_decode_step_match_the_survey_table(tp, batch, want_gb):
    assert step_bytes(LLAMA3_8B, tp, batch, 512) / 1e9 == pytest.approx(want_gb, abs=0.006)


def test_70b_tp8_and_the_planner_defaults():
    assert step_bytes(LLAMA3_70B, 8, 32, 512) / 1e9 == pytest.approx(18.05, abs=0.01)
    # the reference arm and the GPU arm describe the same workload
    a, b = bench.workload_config(1, 32, 512), bench.workload_config(8, 64, 512)
    assert "configs[1]" in a["workload"] and "batch 32" in a["workload"]
    assert "configs[2]" in b["workload"] and "TP=8" in b["workload"] and "batch 64" in b["workload"]
