"""CLIP byte-pair-encoding tokenizer (reference ``language_table/common/clip_tokenizer.py:42-152``, an in-graph TF version of
OpenAI's ``SimpleTokenizer``) in plain Python.

The merge table (``bpe_simple_vocab_16e6.txt.gz``) is not shipped with either repository; pass its path to
``create_vocab``.  Token ids follow CLIP: 256 byte symbols, the same 256 with ``</w>``, one id per merge, then
``<|startoftext|>`` and ``<|endoftext|>``; ``tokenize_text`` pads to 77 ids framed by the start / end tokens.
"""
import gzip
import re
from functools import lru_cache
from typing import Dict, List, Sequence, Tuple

import numpy as np


@lru_cache()
def bytes_to_unicode() -> Dict[int, str]:
    """The reversible byte → printable-unicode map of GPT-2 / CLIP."""
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(ord("¡"), ord("¬") + 1)) + list(range(ord("®"), ord("ÿ") + 1))
    cs, n = bs[:], 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return dict(zip(bs, [chr(c) for c in cs]))


def create_vocab(bpe_path: str = None, merges: Sequence[Tuple[str, str]] = None):
    """Returns ``(token → id, merge → rank)``.  Either a path to the gzipped merge file or an explicit merge list."""
    if merges is None:
        with gzip.open(bpe_path) as f:
            lines = f.read().decode("utf-8").split("\n")
        merges = [tuple(m.split()) for m in lines[1:49152 - 256 - 2 + 1]]
    vocab = list(bytes_to_unicode().values())
    vocab = vocab + [v + "</w>" for v in vocab] + ["".join(m) for m in merges] + ["<|startoftext|>", "<|endoftext|>"]
    return {t: i for i, t in enumerate(vocab)}, {tuple(m): i for i, m in enumerate(merges)}


_PAT = re.compile(r"<\|startoftext\|>|<\|endoftext\|>|'s|'t|'re|'ve|'m|'ll|'d|[^\W\d_]+|\d|[^\s\w]+", re.IGNORECASE | re.UNICODE)


class ClipTokenizer:
    def __init__(self, encoder: Dict[str, int], bpe_ranks: Dict[Tuple[str, str], int]):
        self.encoder, self.bpe_ranks, self.byte_encoder = encoder, bpe_ranks, bytes_to_unicode()
        self.cache = {"<|startoftext|>": "<|startoftext|>", "<|endoftext|>": "<|endoftext|>"}
        self.unk = len(encoder)                        # one out-of-vocabulary bucket, like the reference's lookup table

    def bpe(self, token: str) -> str:
        if token in self.cache:
            return self.cache[token]
        word = tuple(token[:-1]) + (token[-1] + "</w>",)
        while len(word) > 1:
            pairs = {(a, b) for a, b in zip(word, word[1:])}
            best = min(pairs, key=lambda p: self.bpe_ranks.get(p, float("inf")))
            if best not in self.bpe_ranks:
                break
            a, b = best
            out, i = [], 0
            while i < len(word):
                if i < len(word) - 1 and word[i] == a and word[i + 1] == b:
                    out.append(a + b)
                    i += 2
                else:
                    out.append(word[i])
                    i += 1
            word = tuple(out)
        self.cache[token] = " ".join(word)
        return self.cache[token]

    def tokenize(self, text: str) -> List[int]:
        text = re.sub(r"\s+", " ", text.strip()).lower()
        ids = []
        for tok in _PAT.findall(text):
            tok = "".join(self.byte_encoder[b] for b in tok.encode("utf-8"))
            ids.extend(self.encoder.get(t, self.unk) for t in self.bpe(tok).split(" "))
        return ids


def tokenize_text(text, tokenizer: ClipTokenizer, vocab_size: int = 49408, context_length: int = 77) -> np.ndarray:
    """``text``: a string or a list of strings → int64 [n, 77]: start token, BPE ids, end token, zero padding."""
    texts = [text] if isinstance(text, str) else list(text)
    out = np.zeros((len(texts), context_length), dtype=np.int64)
    for i, t in enumerate(texts):
        ids = [vocab_size - 2] + tokenizer.tokenize(t) + [vocab_size - 1]
        out[i, :min(len(ids), context_length)] = ids[:context_length]
    return out
