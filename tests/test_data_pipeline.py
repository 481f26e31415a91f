"""Offline data-prep CLIs end to end on a tiny corpus (reference preprocess_data.py / train_tokenizer.py /
pre_tokenize.py, recipe.sh steps 2-4): parquet -> text JSON -> BPE tokenizer -> token-id JSON the dataset reads."""
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, cwd):
    r = subprocess.run([sys.executable] + args, cwd=cwd, capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, PYTHONPATH=REPO))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


def test_preprocess_tokenizer_pretokenize(tmp_path):
    import pandas as pd
    words = ["the", "king", "queen", "castle", "sword", "glory", "night", "morning", "good", "hello", "world", "this",
             "is", "a", "test", "another", "brave", "old", "man", "duty"]
    docs = [" ".join(words[(i * 7 + j * 3) % len(words)] for j in range(5 + i % 30)) for i in range(400)]
    docs.append("x" * 5000)           # longer than 2000 chars: must be filtered out
    docs += ["good morning", "hello world", "this is a test", "this is another test"] * 5
    pq = tmp_path / "raw.parquet"
    pd.DataFrame({"text": docs, "other": range(len(docs))}).to_parquet(pq)
    out = _run([os.path.join(REPO, "preprocess_data.py"), str(pq), str(tmp_path / "data.json"), "--seed", "0"],
               str(tmp_path))
    assert "train:" in out and "validation:" in out and "documents" in out
    data = json.load(open(tmp_path / "data.json"))
    assert set(data) == {"train", "validation"} and len(data["train"]) + len(data["validation"]) == len(docs) - 1
    assert len(data["validation"]) >= 1 and all(len(d) <= 2000 for d in data["train"])
    # bare output filename (reference defect D10: os.makedirs('') crashed)
    _run([os.path.join(REPO, "train_tokenizer.py"), "-d", "data.json", "-v", "300", "-o", "tok.json"], str(tmp_path))
    from tokenizers import Tokenizer
    tok = Tokenizer.from_file(str(tmp_path / "tok.json"))
    assert [tok.token_to_id(t) for t in ("<BOS>", "<EOS>", "<UNK>")] == [0, 1, 2] and tok.get_vocab_size() <= 300
    out = _run([os.path.join(REPO, "pre_tokenize.py"), "-i", "data.json", "-o", "tokens.json", "-t", "tok.json"],
               str(tmp_path))
    assert "train:" in out and "mean" in out and "longest" in out
    tk = json.load(open(tmp_path / "tokens.json"))
    assert set(tk) == {"train", "validation", "special_ids", "vocab_size"}
    assert tk["special_ids"] == {"<BOS>": 0, "<EOS>": 1, "<UNK>": 2} and tk["vocab_size"] == tok.get_vocab_size()
    sys.path.insert(0, REPO)
    from dataset import get_dataloader
    batch = next(iter(get_dataloader(str(tmp_path / "tokens.json"), 4, -1, "train", maxlen=64, shuffle=False)))
    assert batch["input_ids"][0, 0].item() == 0 and batch["input_ids"].max().item() < tk["vocab_size"]
