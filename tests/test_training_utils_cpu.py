"""Training-driver utilities: schedules, param groups, TensorBoard logger, datasets (flat token file, HF datasets directory,
instruction data, BERT shards) and the device prefetcher (reference ``examples/training/llama/training_utils.py``)."""
import os, tempfile, json, numpy as np, torch
from types import SimpleNamespace
from neuronx_distributed_llama3_2_b200.utils import training_utils as tu, data as D


def test_training_utils_and_datasets():
    # schedules / groups
    m = torch.nn.Sequential(torch.nn.Linear(4, 4), torch.nn.LayerNorm(4))
    groups = tu.get_param_groups_by_weight_decay(m, 0.1)
    assert len(groups[0]["params"]) == 1 and len(groups[1]["params"]) == 3
    opt = torch.optim.AdamW(groups, lr=1.0)
    sch = tu.get_learning_rate_scheduler(opt, SimpleNamespace(max_steps=10, min_lr=0.1, warmup_steps=2, constant_steps=2))
    lrs = []
    for _ in range(12):
        lrs.append(opt.param_groups[0]["lr"]); opt.step(); sch.step()
    assert lrs[0] < lrs[1] < lrs[2] and abs(lrs[-1] - 0.1) < 1e-9 and max(lrs) <= 1.0, lrs
    assert tu.create_partition(8, 4) == ["model.layers.1", "model.layers.3", "model.layers.5"]
    sin, cos = tu.get_sin_cos_matrix(SimpleNamespace(hidden_size=64, num_attention_heads=4, max_position_embeddings=16, rope_theta=1e4))
    assert sin.shape == (16, 8) and torch.allclose(sin ** 2 + cos ** 2, torch.ones(16, 8))
    assert tu.get_mixed_precision_config(1)["use_fp32_grad_acc"] and tu.get_dtype(m) == "torch.float32"
    with tempfile.TemporaryDirectory() as d:
        lg = tu.TrainingLogger(SimpleNamespace(output_dir=d, lr=1e-3, batch_size=2, grad_accum_usteps=1, warmup_steps=1, max_steps=3), 1, "torch.float32")
        tu.print_logs(torch.tensor(2.5), torch.tensor(1.25), None, 7.0, lg, 1, 1e-3, torch.zeros(2, 4), 0.0)
        lg.close()
        ev = [os.path.join(lg.log_dir, f) for f in os.listdir(lg.log_dir)][0]
        sc = tu.load_events(ev)
        assert abs(sc["step loss"][0].value - 2.5) < 1e-6 and "grad-norm" in sc
        # flat token file
        toks = np.arange(1000, dtype=np.uint16)
        np.save(os.path.join(d, "tok.npy"), toks)
        dl, _ = D.create_llama_pretraining_dataset(os.path.join(d, "tok.npy"), 2, 2, 1, 0, seq_len=100)
        b = next(iter(dl))
        assert b["input_ids"].shape == (2, 100) and b["input_ids"][0, 0].item() == 100    # dp rank 1 starts at window 1
        assert torch.equal(b["input_ids"], b["labels"])
        pf = D.DevicePrefetcher(dl, "cpu")
        got = [next(pf)["input_ids"][0, 0].item() for _ in range(5)]                      # loops over epochs
        assert got[:2] == [100, 500] and len(got) == 5, got
        # HF datasets directory
        import datasets
        ds = datasets.Dataset.from_dict({"input_ids": D.pack_dataset([[1, 2, 3], list(range(10, 20))], 4)})
        ds.save_to_disk(os.path.join(d, "hf"))
        dl2, _ = D.create_llama_pretraining_dataset(os.path.join(d, "hf"), 1, 1, 0, 0)
        rows = [b["input_ids"][0].tolist() for b in dl2]
        assert rows == [[1, 2, 3, 10], [11, 12, 13, 14], [15, 16, 17, 18]], rows
        # instruction data
        with open(os.path.join(d, "train.jsonl"), "w") as f:
            for i in range(8):
                f.write(json.dumps({"instruction": f"add {i}", "input": "", "output": f"{i + 1}"}) + "\n")
        tr, ev2 = D.create_instruction_based_dataset(d, 2, 1, 0, 0, tokenizer=lambda s: [ord(c) % 250 + 1 for c in s], seq_len=256)
        b = next(iter(tr))
        n_prompt = (b["labels"][0] == -100).sum().item()
        assert b["input_ids"].shape == (2, 256) and 0 < (b["labels"][0] >= 0).sum().item() <= 2 and ev2 is None
        assert b["attention_mask"][0].sum().item() < 256 and n_prompt > 100
        # BERT shard
        z = dict(input_ids=np.random.randint(1, 100, (3, 16)), input_mask=np.ones((3, 16), dtype=np.int64), segment_ids=np.zeros((3, 16), dtype=np.int64),
                 masked_lm_positions=np.array([[2, 5, 0, 0]] * 3), masked_lm_ids=np.array([[7, 9, 0, 0]] * 3), next_sentence_labels=np.array([0, 1, 0]))
        np.savez(os.path.join(d, "shard.npz"), **z)
        bd = D.BertPretrainingDataset(os.path.join(d, "shard.npz"), max_pred_length=4)
        ids, seg, mask, labels, nsl = bd[1]
        assert labels[2] == 7 and labels[5] == 9 and (labels >= 0).sum() == 2 and nsl == 1
