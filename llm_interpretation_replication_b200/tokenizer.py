"""Deterministic offline tokenizer shared by the oracle and the CUDA path.

No tokenizer files exist offline, so both sides use the same 4-byte-chunk hash: the UTF-8
string is cut into consecutive 4-byte chunks (the trailing partial chunk is kept) and
``id = 32 + crc32(chunk) mod (V - 64)``.  The class exposes exactly the members the
reference scorer touches (reference ``analysis/run_base_vs_instruct_100q.py:329-345,365-368``):
``__call__(text, return_tensors=, add_special_tokens=, padding=)`` -> object with
``.input_ids`` and ``.to()``, ``.encode``, ``.decode(skip_special_tokens=)``,
``.pad_token[_id]``, ``.eos_token[_id]``.
"""
from __future__ import annotations

import zlib
from typing import List, Sequence


class _Encoding(dict):
    def __init__(self, input_ids):
        super().__init__(input_ids=input_ids)
        self.input_ids = input_ids

    def to(self, device):
        if hasattr(self.input_ids, "to"):
            return _Encoding(self.input_ids.to(device))
        return self


class HashTokenizer:
    CHUNK = 4

    def __init__(self, vocab: int, eos_token_id: int = 11, bos_token_id: int | None = None,
                 prepend_bos: bool = False, overrides: dict | None = None, append_eos: bool = False,
                 pad_token_id: int | None = None):
        assert vocab > 64 + 8
        self.vocab_size = vocab
        self.eos_token_id = eos_token_id
        self.eos_token = "<|endoftext|>"
        self.pad_token_id = pad_token_id
        self.pad_token = None if pad_token_id is None else "<pad>"
        self.append_eos = append_eos     # emulates the T5 tokenizers, which close every encoded string with </s>
        self.bos_token_id = bos_token_id
        self.prepend_bos = prepend_bos   # emulates BOS-prepending tokenizers
        self.overrides = dict(overrides or {})   # whole-string -> single id (tests pin candidate ids)

    # -- core hash -------------------------------------------------------------------
    def _ids(self, text: str) -> List[int]:
        if text in self.overrides:
            return [int(self.overrides[text])]
        b = text.encode("utf-8")
        out = []
        for i in range(0, len(b), self.CHUNK):
            out.append(32 + zlib.crc32(b[i:i + self.CHUNK]) % (self.vocab_size - 64))
        return out

    def encode(self, text: str, add_special_tokens: bool = True) -> List[int]:
        ids = self._ids(text)
        if add_special_tokens and self.prepend_bos and self.bos_token_id is not None:
            ids = [self.bos_token_id] + ids
        if add_special_tokens and self.append_eos:
            ids = ids + [self.eos_token_id]
        return ids

    def __call__(self, text, return_tensors=None, add_special_tokens: bool = True, padding=False):
        ids = self.encode(text, add_special_tokens=add_special_tokens)
        if return_tensors == "pt":
            import torch
            return _Encoding(torch.tensor([ids], dtype=torch.long))
        return _Encoding(ids)

    def decode(self, ids: Sequence[int], skip_special_tokens: bool = False) -> str:
        toks = []
        for t in ids:
            t = int(t)
            if skip_special_tokens and (t == self.eos_token_id or (self.pad_token_id is not None and t == self.pad_token_id)):
                continue
            toks.append(f"<{t}>")
        return "".join(toks)

    def apply_pad_default(self):
        # reference run_base_vs_instruct_100q.py:422-423
        if self.pad_token is None:
            self.pad_token = self.eos_token
            self.pad_token_id = self.eos_token_id
        return self
