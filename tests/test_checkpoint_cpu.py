"""Checkpoint layout under a mocked topology, retention / async / xser paths, S3 storage against an in-memory client
(reference ``test/unit_test/checkpoint/test_{checkpoint,checkpoint_storage}.py``: one process believes it is
``tp_rank=1, pp_rank=1`` of a TP8 x PP8 world and runs the real save / load code)."""
from __future__ import annotations

import datetime
import io
import os

import pytest
import torch

import neuronx_distributed_llama3_2_b200 as nxd
from neuronx_distributed_llama3_2_b200.parallel import parallel_state as ps
from neuronx_distributed_llama3_2_b200.trainer import checkpoint as ckpt
from neuronx_distributed_llama3_2_b200.trainer.checkpoint_storage import (FilesysCheckpointStorage, S3CheckpointStorage,
                                                                        create_checkpoint_storage)


class _Net(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(8, 16)
        self.b = torch.nn.Linear(16, 4)

    def forward(self, x):
        return self.b(torch.relu(self.a(x)))


@pytest.fixture
def mocked_topology(monkeypatch):
    """tp_rank 1 of 8, pp_rank 1 of 8, dp_rank 0 -- through the getters the checkpoint code calls."""
    monkeypatch.setattr(ckpt, "_mp", lambda: True)
    monkeypatch.setattr(ckpt, "_ranks", lambda: (0, 1, 1, 0, 1))
    monkeypatch.setattr(ckpt, "_rendezvous", lambda name: None)
    yield


def _files(root):
    out = []
    for d, _, fs in os.walk(root):
        for f in fs:
            out.append(os.path.relpath(os.path.join(d, f), root))
    return sorted(out)


@pytest.mark.parametrize("use_xser", [False, True])
def test_layout_and_round_trip_under_mocked_topology(tmp_path, mocked_topology, use_xser):
    torch.manual_seed(0)
    model, opt = _Net(), None
    opt = torch.optim.AdamW(model.parameters(), lr=1e-2)
    model(torch.randn(4, 8)).sum().backward()
    opt.step()
    sched = torch.optim.lr_scheduler.StepLR(opt, 3)
    root = str(tmp_path / "ckpts")
    nxd.save_checkpoint(root, "unittest", model=model, optimizer=opt, scheduler=sched, user_content={"total_steps": 7},
                        use_xser=use_xser)
    nxd.finalize_checkpoint()
    files = _files(root)
    assert "unittest/done" in files and "unittest/checkpoint" in files
    assert "unittest/model/dp_rank_00_tp_rank_01_pp_rank_01.pt" in files
    assert "unittest/optim/dp_rank_00_tp_rank_01_pp_rank_01.pt" in files
    assert any(f.startswith("unittest/scheduler") for f in files) and any(f.startswith("unittest/user_content") for f in files)
    assert any(".tensors/" in f for f in files) == use_xser
    assert nxd.has_checkpoint(root)
    want = {k: v.clone() for k, v in model.state_dict().items()}
    want_opt = opt.state_dict()
    model2 = _Net()
    opt2 = torch.optim.AdamW(model2.parameters(), lr=1.0)
    user = nxd.load_checkpoint(root, tag=None, model=model2, optimizer=opt2)      # tag None = newest completed
    assert user == {"total_steps": 7}
    for k, v in model2.state_dict().items():
        assert torch.equal(v, want[k]), k
    got_opt = opt2.state_dict()
    assert got_opt["param_groups"][0]["lr"] == want_opt["param_groups"][0]["lr"]
    for i, st in want_opt["state"].items():
        assert torch.equal(got_opt["state"][i]["exp_avg"], st["exp_avg"])


def test_retention_async_and_incomplete_tags(tmp_path, mocked_topology):
    model = _Net()
    root = str(tmp_path / "ckpts")
    for step in (1, 2, 3, 4):
        nxd.save_checkpoint(root, f"step_{step}", model=model, num_kept_ckpts=2, async_save=True)
    nxd.finalize_checkpoint()
    store = FilesysCheckpointStorage(root)
    assert store.list_completed_checkpoint_tags() == ["step_3", "step_4"]
    # a tag without its `done` marker (a save that died half way) is never picked up
    os.makedirs(os.path.join(root, "step_5", "model"))
    open(os.path.join(root, "step_5", "checkpoint"), "w").close()
    assert store.list_completed_checkpoint_tags() == ["step_3", "step_4"]
    assert "step_5" in store.list_checkpoint_tags()
    model2 = _Net()
    nxd.load_checkpoint(root, model=model2)
    assert torch.equal(model2.a.weight, model.a.weight)
    assert not nxd.has_checkpoint(str(tmp_path / "nothing_here"))


class _FakeS3:
    """In-memory stand-in for the boto3 client calls the storage class makes."""

    def __init__(self):
        self.objects = {}
        self.calls = []
        self.fail_next = 0
        self._uploads = {}

    def _maybe_fail(self):
        if self.fail_next > 0:
            self.fail_next -= 1
            raise RuntimeError("An error occurred (SlowDown) when calling the PutObject operation")

    def put_object(self, Bucket, Key, Body):
        self.calls.append(("put", Key))
        self._maybe_fail()
        self.objects[Key] = (bytes(Body), datetime.datetime.now())

    def get_object(self, Bucket, Key):
        return {"Body": io.BytesIO(self.objects[Key][0])}

    def list_objects_v2(self, Bucket, Prefix, Delimiter=None, ContinuationToken=None):
        keys = sorted(k for k in self.objects if k.startswith(Prefix))
        if not Delimiter:
            return {"Contents": [{"Key": k, "LastModified": self.objects[k][1]} for k in keys], "IsTruncated": False}
        contents, prefixes = [], set()
        for k in keys:
            rest = k[len(Prefix):]
            if Delimiter in rest:
                prefixes.add(Prefix + rest.split(Delimiter)[0] + Delimiter)
            else:
                contents.append({"Key": k, "LastModified": self.objects[k][1]})
        return {"Contents": contents, "CommonPrefixes": [{"Prefix": p} for p in sorted(prefixes)], "IsTruncated": False}

    def delete_object(self, Bucket, Key):
        self.objects.pop(Key, None)

    def delete_objects(self, Bucket, Delete):
        for o in Delete["Objects"]:
            self.objects.pop(o["Key"], None)

    def create_multipart_upload(self, Bucket, Key):
        self._uploads[Key] = {}
        return {"UploadId": Key}

    def upload_part(self, Bucket, Key, PartNumber, UploadId, Body):
        self._uploads[Key][PartNumber] = bytes(Body)
        return {"ETag": f"etag{PartNumber}"}

    def complete_multipart_upload(self, Bucket, Key, UploadId, MultipartUpload):
        parts = [self._uploads[Key][p["PartNumber"]] for p in MultipartUpload["Parts"]]
        self.objects[Key] = (b"".join(parts), datetime.datetime.now())

    def abort_multipart_upload(self, Bucket, Key, UploadId):
        self._uploads.pop(Key, None)


def test_s3_storage_with_in_memory_client(monkeypatch):
    assert isinstance(create_checkpoint_storage("/tmp/x"), FilesysCheckpointStorage)
    assert isinstance(create_checkpoint_storage("s3://bucket/run1"), S3CheckpointStorage)
    with pytest.raises(ValueError):
        S3CheckpointStorage("/not/s3")
    fake = _FakeS3()
    monkeypatch.setattr("time.sleep", lambda s: None)
    st = S3CheckpointStorage("s3://bucket/run1", client=fake)
    st.save_text("1", "step_1/checkpoint")
    st.save_object({"w": torch.arange(6)}, "step_1/model/dp_rank_00_tp_rank_00_pp_rank_00.pt")
    fake.fail_next = 2                                       # throttled twice, then succeeds
    st.save_text("1", "step_1/done")
    assert [c for c in fake.calls if c[1].endswith("step_1/done")].__len__() == 3
    st.save_text("1", "step_2/checkpoint")                   # incomplete
    assert st.file_exists("step_1/done") and not st.file_exists("step_2/done")
    assert st.dir_exists("step_1") and not st.dir_exists("step_9")
    assert st.list_completed_checkpoint_tags() == ["step_1"]
    assert sorted(st.list_checkpoint_tags()) == ["step_1", "step_2"]
    back = st.load_object("step_1/model/dp_rank_00_tp_rank_00_pp_rank_00.pt")
    assert torch.equal(back["w"], torch.arange(6))
    assert st.find_files("step_1", r"tp_rank_00") == ["step_1/model/dp_rank_00_tp_rank_00_pp_rank_00.pt"]
    # multipart path: lower the part size instead of allocating 64 MiB
    monkeypatch.setattr(S3CheckpointStorage, "PART_SIZE", 1000)
    st.save_object({"w": torch.arange(2000)}, "step_1/big.pt")
    assert torch.equal(st.load_object("step_1/big.pt")["w"], torch.arange(2000))
    assert not st.is_checkpoint_xser("step_1")


def test_nxd_config_defaults_and_wrappers():
    cfg = nxd.neuronx_distributed_config()
    assert cfg["tensor_parallel_size"] == 1 and cfg["pipeline_parallel_size"] == 1
    assert cfg["optimizer_config"]["zero_one_enabled"] is False and cfg["optimizer_config"]["grad_clipping"] is True
    assert "mixed_precision_config" in cfg and "activation_checkpoint_config" in cfg and "model_init_config" in cfg
    cfg = nxd.neuronx_distributed_config(optimizer_config={"zero_one_enabled": True, "max_grad_norm": 0.5})
    assert cfg["optimizer_config"]["max_grad_norm"] == 0.5 and cfg["optimizer_config"]["grad_clipping"] is True
    model = nxd.initialize_parallel_model(nxd.neuronx_distributed_config(), _Net)
    opt = nxd.initialize_parallel_optimizer(nxd.neuronx_distributed_config(), torch.optim.SGD, model.parameters(), lr=0.1)
    assert model.local_modules()[0] is model.original_module() or len(model.local_modules()) == 1
    assert {n for n, _ in model.named_parameters()} == {"a.weight", "a.bias", "b.weight", "b.bias"}
    class _WithLoss(_Net):
        def forward(self, x):
            return super().forward(x).pow(2).mean()

    model = nxd.initialize_parallel_model(nxd.neuronx_distributed_config(), _WithLoss)
    opt = nxd.initialize_parallel_optimizer(nxd.neuronx_distributed_config(), torch.optim.SGD, model.parameters(), lr=0.1)
    before = model.original_module().a.weight.clone()
    loss = model.run_train(torch.randn(3, 8))             # forward + backward, returns the loss (reference semantics)
    assert loss.dim() == 0 and model.original_module().a.weight.grad is not None
    opt.step()
    assert float(opt.grad_norm) > 0 and not torch.equal(before, model.original_module().a.weight)
    assert model.run_eval(torch.randn(3, 8)).dim() == 0
    assert ps.get_tensor_model_parallel_size() == 1
