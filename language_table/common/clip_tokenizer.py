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

    @staticmethod
    def clean(text: str) -> str:
        """Strip, collapse white space, lower-case (``clip_tokenizer.py:72-84``; like there, no HTML un-escaping)."""
        return re.sub(r"\s+", " ", text.strip()).lower()

    def tokenize(self, text: str) -> List[int]:
        return self.tokenize_with_offsets(text)[0]

    def tokenize_with_offsets(self, text: str):
        """``(ids, begin, end)``: one entry per BPE piece; ``begin`` / ``end`` are byte offsets (end exclusive) into the
        UTF-8 encoding of the cleaned text (``clean``), so ``clean(text).encode()[b:e]`` is the piece's surface form.
        The reference's method of the same name (``clip_tokenizer.py:71-110``) returns ragged tensors with the same three
        parts; its end offsets also count the four characters of the ``</w>`` marker, which are not text."""
        text = self.clean(text)
        ids, begin, end = [], [], []
        char_to_byte = np.cumsum([0] + [len(c.encode("utf-8")) for c in text])
        for m in _PAT.finditer(text):
            at = int(char_to_byte[m.start()])
            tok = "".join(self.byte_encoder[b] for b in m.group().encode("utf-8"))      # one symbol per byte
            for piece in self.bpe(tok).split(" "):
                n = len(piece[:-4] if piece.endswith("</w>") else piece)
                ids.append(self.encoder.get(piece, self.unk))
                begin.append(at)
                end.append(at + n)
                at += n
        return ids, begin, end


def tokenize_text(text, tokenizer: ClipTokenizer, vocab_size: int = 49408, context_length: int = 77) -> np.ndarray:
    """``text``: a string or a list of strings → int64 [n, 77]: start token, BPE ids, end token, zero padding."""
    texts = [text] if isinstance(text, str) else list(text)
    out = np.zeros((len(texts), context_length), dtype=np.int64)
    for i, t in enumerate(texts):
        ids = [vocab_size - 2] + tokenizer.tokenize(t) + [vocab_size - 1]
        out[i, :min(len(ids), context_length)] = ids[:context_length]
    return out
