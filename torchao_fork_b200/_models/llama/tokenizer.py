"""Tokenizers for the gpt-fast Llama: sentencepiece (Llama-2 family) and tiktoken BPE (Llama-3 family).

Counterpart of /root/reference/torchao/_models/llama/tokenizer.py:13-126 (what `generate.py` / `eval.py`
put in front of the hot path).  Same surface: `encode(text) -> list[int]`, `decode(ids) -> str`, `bos_id()`, `eos_id()`,
and `get_tokenizer(path, model_name)` picking the tiktoken one when the model name says "Llama-3".

The Llama-3 vocabulary is the BPE rank file plus 256 specials numbered right after the base tokens; the order of the first
ten is fixed by the published model (begin / end of text, four reserved, the two header markers, one reserved, end of
turn), the rest are `<|reserved_special_token_N|>`.
"""
import os
from pathlib import Path
from typing import Dict, List

# split pattern of the Llama-3 tokenizer (tokenizer.py:61)
_LLAMA3_SPLIT = (r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+"
                 r"|\s+(?!\S)|\s+")
_LLAMA3_N_SPECIAL = 256
_LLAMA3_NAMED = {0: "<|begin_of_text|>", 1: "<|end_of_text|>", 6: "<|start_header_id|>", 7: "<|end_header_id|>",
                 9: "<|eot_id|>"}


def llama3_special_tokens(num_base_tokens: int) -> Dict[str, int]:
    """name -> id of the 256 specials: the named ones at their slots, `reserved_special_token_N` filling the gaps in order."""
    out, reserved = {}, 0
    for slot in range(_LLAMA3_N_SPECIAL):
        name = _LLAMA3_NAMED.get(slot)
        if name is None:
            name = f"<|reserved_special_token_{reserved}|>"
            reserved += 1
        out[name] = num_base_tokens + slot
    return out


class TokenizerInterface:
    """What generate / eval call.  Subclasses set the `_enc` / `_dec` callables and the two ids;
    a bare instance raises, as the reference's abstract base does."""

    def __init__(self, model_path):
        self.model_path = model_path
        self._enc = self._dec = None
        self.bos_token_id = self.eos_token_id = None

    def _need(self, what):
        if what is None:
            raise NotImplementedError("This method should be overridden by subclasses.")
        return what

    def encode(self, text: str) -> List[int]:
        return self._need(self._enc)(text)

    def decode(self, tokens: List[int]) -> str:
        return self._need(self._dec)(tokens)

    def bos_id(self) -> int:
        return self._need(self.bos_token_id)

    def eos_id(self) -> int:
        return self._need(self.eos_token_id)


class SentencePieceWrapper(TokenizerInterface):
    def __init__(self, model_path):
        from sentencepiece import SentencePieceProcessor
        super().__init__(model_path)
        self.processor = sp = SentencePieceProcessor(str(model_path))
        self._enc, self._dec = sp.EncodeAsIds, sp.DecodeIds
        self.bos_token_id, self.eos_token_id = sp.bos_id(), sp.eos_id()


class TiktokenWrapper(TokenizerInterface):
    def __init__(self, model_path):
        import tiktoken
        from tiktoken.load import load_tiktoken_bpe
        super().__init__(model_path)
        if not os.path.isfile(model_path):
            raise FileNotFoundError(str(model_path))
        ranks = load_tiktoken_bpe(str(model_path))
        self.special_tokens = llama3_special_tokens(len(ranks))
        self.model = tiktoken.Encoding(name=Path(model_path).name, pat_str=_LLAMA3_SPLIT, mergeable_ranks=ranks,
                                       special_tokens=self.special_tokens)
        self._enc, self._dec = self.model.encode, self.model.decode  # encode: specials in the text are an error, as upstream
        self.bos_token_id = self.special_tokens["<|begin_of_text|>"]
        self.eos_token_id = self.special_tokens["<|end_of_text|>"]


def get_tokenizer(tokenizer_model_path, model_name) -> TokenizerInterface:
    """reference: tokenizer.py:113-126"""
    if "Llama-3" in str(model_name):
        return TiktokenWrapper(tokenizer_model_path)
    return SentencePieceWrapper(tokenizer_model_path)
