"""Tokenizer wrappers (reference: torchao/_models/llama/tokenizer.py).  No tokenizer model ships offline, so the tiktoken
one is exercised on a byte-level rank file written here, the sentencepiece one on a model trained here from a few lines."""
import base64

import pytest

from torchao_fork_b200._models.llama import tokenizer as tk


def test_llama3_special_token_table():
    sp = tk.llama3_special_tokens(128000)
    assert len(sp) == 256 and sorted(sp.values()) == list(range(128000, 128256))
    assert sp["<|begin_of_text|>"] == 128000 and sp["<|end_of_text|>"] == 128001
    assert sp["<|reserved_special_token_0|>"] == 128002 and sp["<|reserved_special_token_3|>"] == 128005
    assert sp["<|start_header_id|>"] == 128006 and sp["<|end_header_id|>"] == 128007
    assert sp["<|reserved_special_token_4|>"] == 128008 and sp["<|eot_id|>"] == 128009
    assert sp["<|reserved_special_token_5|>"] == 128010 and sp["<|reserved_special_token_250|>"] == 128255


def test_tiktoken_wrapper_roundtrip(tmp_path):
    pytest.importorskip("tiktoken")
    path = tmp_path / "tokenizer.model"
    merges = [bytes([b]) for b in range(256)] + [b"he", b"ll", b"hell", b"hello", b" w", b"or"]
    path.write_text("".join(f"{base64.b64encode(tok).decode()} {rank}\n" for rank, tok in enumerate(merges)))
    t = tk.get_tokenizer(path, "checkpoints/meta-llama/Meta-Llama-3-8B/model.pth")
    assert isinstance(t, tk.TiktokenWrapper)
    assert t.bos_id() == len(merges) and t.eos_id() == len(merges) + 1 == t.eos_token_id
    text = "hello world, 12345 café\n"
    ids = t.encode(text)
    assert ids[0] == 256 + 3 and t.decode(ids) == text  # "hello" is one token
    assert all(i < len(merges) for i in ids)  # ordinary text never yields a special
    assert t.decode([t.bos_id()]) == "<|begin_of_text|>"
    with pytest.raises(FileNotFoundError):
        tk.TiktokenWrapper(tmp_path / "missing.model")


def test_sentencepiece_wrapper_roundtrip(tmp_path):
    spm = pytest.importorskip("sentencepiece")
    corpus = tmp_path / "corpus.txt"
    corpus.write_text("\n".join(["the quick brown fox jumps over the lazy dog", "hello my name is llama",
                                 "weights are packed four bits at a time"] * 20))
    try:
        spm.SentencePieceTrainer.train(input=str(corpus), model_prefix=str(tmp_path / "tokenizer"), vocab_size=60,
                                       model_type="bpe", bos_id=1, eos_id=2, unk_id=0, minloglevel=2)
    except Exception as e:  # a build without the trainer
        pytest.skip(f"sentencepiece trainer unavailable: {e}")
    t = tk.get_tokenizer(tmp_path / "tokenizer.model", "Llama-2-7b-chat-hf")
    assert isinstance(t, tk.SentencePieceWrapper)
    assert (t.bos_id(), t.eos_id()) == (1, 2) == (t.bos_token_id, t.eos_token_id)
    ids = t.encode("hello my name is llama")
    assert ids and t.decode(ids) == "hello my name is llama"


def test_interface_is_abstract():
    base = tk.TokenizerInterface("x")
    for call in (lambda: base.encode("a"), lambda: base.decode([1]), base.bos_id, base.eos_id):
        with pytest.raises(NotImplementedError):
            call()
