"""Host-side pieces of the payload that need no GPU: dataset ingest from the dataset mount,
deterministic batching, and that the metrics file has the shape the controller's
S3Handler.get_metrics (reference app/utils/S3Handler.py:252-286) turns into records."""
import json
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
from b200ft import train  # noqa: E402


def test_token_stream_formats(tmp_path):
    assert train.load_token_stream(str(tmp_path)) is None          # empty mount -> synthetic fallback
    ids = np.arange(1000, dtype=np.uint16)
    d = tmp_path / "bin"; d.mkdir()
    ids.tofile(d / "tokens.bin")
    s = train.load_token_stream(str(d))
    assert s.dtype == torch.int64 and torch.equal(s, torch.arange(1000))
    d = tmp_path / "pt"; d.mkdir()
    torch.save(torch.arange(50, dtype=torch.int32).view(5, 10), d / "tokens.pt")
    assert torch.equal(train.load_token_stream(str(d)), torch.arange(50))
    d = tmp_path / "jsonl"; d.mkdir()
    with open(d / "a.jsonl", "w") as f:
        f.write(json.dumps({"input_ids": [1, 2, 3]}) + "\n\n" + json.dumps({"input_ids": [4, 5]}) + "\n")
    assert train.load_token_stream(str(d)).tolist() == [1, 2, 3, 4, 5]


def test_batches_are_deterministic_disjoint_across_ranks_and_wrap():
    stream = torch.arange(10_000)
    b0 = train.batch_from_stream(stream, 3, 0, 2, 4, 16, vocab=50257)
    b1 = train.batch_from_stream(stream, 3, 1, 2, 4, 16, vocab=50257)
    assert b0.shape == (4, 16) and torch.equal(b0, train.batch_from_stream(stream, 3, 0, 2, 4, 16, vocab=50257))
    assert set(b0.flatten().tolist()).isdisjoint(b1.flatten().tolist())
    assert int(b1[0, 0]) == int(b0[0, 0]) + 64                      # consecutive spans, ranks interleaved
    short = torch.arange(40)
    w = train.batch_from_stream(short, 0, 0, 1, 4, 16, vocab=50257)  # stream shorter than one batch: wraps
    assert w.flatten().tolist() == [i % 40 for i in range(64)]


def test_out_of_range_token_ids_stop_the_job_instead_of_being_clamped():
    train.check_token_range(torch.arange(100), vocab=100)
    with pytest.raises(ValueError, match="vocabulary"):
        train.check_token_range(torch.tensor([1, 2, 100]), vocab=100)
    with pytest.raises(ValueError, match="vocabulary"):
        train.check_token_range(torch.tensor([-1, 2]), vocab=100)
    with pytest.raises(ValueError):
        train.check_token_range(torch.tensor([], dtype=torch.int64), vocab=100)


def test_metrics_csv_is_what_the_controller_parses(tmp_path):
    pd = pytest.importorskip("pandas")
    import csv
    path = tmp_path / "metrics.csv"
    with open(path, "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(["epoch", "step", "loss", "lr", "tokens_per_sec", "elapsed_s"])
        w.writerow([1, 10, "10.812345", 1e-4, "1234567.0", "0.31"])
        w.writerow([1, 20, "10.700000", 1e-4, "1250000.5", "0.57"])
    # what S3Handler.get_metrics does with the newest *metrics*.csv object
    assert "metrics" in path.name and path.name.endswith(".csv")
    records = pd.read_csv(path).fillna(0).to_dict("records")
    assert len(records) == 2 and records[1]["step"] == 20 and abs(records[0]["loss"] - 10.812345) < 1e-9


def test_parser_defaults_match_plugin_defaults():
    ns = train.build_parser().parse_args([])
    assert ns.model == "gpt2-124m" and ns.lora_r == 8 and ns.seq_len == 1024 and ns.batch_size == 32
    assert ns.dataset_path == "/data/dataset" and ns.checkpoint_path == "/data/artifacts" and ns.save_interval == 0
