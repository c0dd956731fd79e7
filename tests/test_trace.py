"""Tracing / step timing helpers: CPU behaviour of StepTimer, TraceWriter and
the trainer's per-epoch device-time report."""
import json
import os

import torch

from ddp_b200.utils.trace import StepTimer, TraceWriter, nvtx_mark, nvtx_range


def test_nvtx_range_is_noop_without_cuda():
    with nvtx_range("outer"):
        with nvtx_range("inner"):
            nvtx_mark("here")


def test_step_timer_cpu_counts_and_resets():
    t = StepTimer("cpu", capacity=4)
    for _ in range(10):           # more steps than the ring holds
        t.begin()
        sum(range(1000))
        t.end()
    s = t.summary()
    assert s["steps"] == 10 and 0 < s["ms_mean"] <= s["ms_max"]
    assert t.summary()["steps"] == 0


def test_trace_writer_disabled_by_default(tmp_path, monkeypatch):
    monkeypatch.delenv("DDPB_TRACE", raising=False)
    tw = TraceWriter()
    assert not tw.enabled
    with tw.span("x"):
        pass
    tw.counter("loss", loss=1.0)
    tw.close()
    assert not list(tmp_path.iterdir())


def test_trace_writer_writes_chrome_trace(tmp_path):
    tw = TraceWriter(str(tmp_path / "tr"), rank=3)
    with tw.span("epoch 0", batches=2):
        tw.counter("loss", loss=2.3)
    tw.close()
    data = json.load(open(tmp_path / "tr.rank3.json"))
    names = [e["name"] for e in data["traceEvents"]]
    assert names == ["loss", "epoch 0"]
    span = data["traceEvents"][1]
    assert span["ph"] == "X" and span["dur"] >= 0 and span["pid"] == 3 and span["args"] == {"batches": 2}


def test_trainer_reports_device_step_time_and_trace(tmp_path, monkeypatch):
    from ddp_b200.trainer import ddp_train
    from ddp_b200.utils.data import synthetic_mnist
    monkeypatch.setenv("DDPB_TRACE", str(tmp_path / "trace"))
    monkeypatch.setenv("MASTER_ADDR", "127.0.0.1")
    monkeypatch.setenv("MASTER_PORT", "29641")
    stats = {}
    ddp_train(0, 1, 1, 16, data_root=str(tmp_path / "data"), ckpt_dir=str(tmp_path / "ck"),
              dataset=synthetic_mnist(64, seed=0), backend="gloo", max_batches=3, stats=stats)
    assert len(stats["device_ms_per_step"]) == 1 and stats["device_ms_per_step"][0] > 0
    data = json.load(open(tmp_path / "trace.rank0.json"))
    names = {e["name"] for e in data["traceEvents"]}
    assert {"epoch 0", "loss", "throughput"} <= names
    assert os.path.exists(tmp_path / "ck" / "epoch_0.pt")
