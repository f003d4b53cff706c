"""Pure-Python equivalents of the torchtext pieces the reference uses (torchtext is not installable offline):
``get_tokenizer('basic_english')``, ``build_vocab_from_iterator(min_freq, specials, special_first)``, ``Vocab``
(``set_default_index``, ``__getitem__``, ``__len__``), and the transforms ``VocabTransform``, ``AddToken``, ``Truncate``,
``ToTensor(padding_value)``, ``PadTransform(max_length, pad_value)``, ``Sequential``
(distributed_lstm.py:75-107; pytorch_machine_translator.py:20-98).  ``get_tokenizer('spacy', ...)`` falls back to a
punctuation-splitting tokenizer (spacy models are not on the box).
"""
from __future__ import annotations

import re
from collections import Counter

import torch

_BASIC_PATTERNS = [(r"\'", " '  "), (r"\"", ""), (r"\.", " . "), (r"<br \/>", " "), (r",", " , "), (r"\(", " ( "),
                   (r"\)", " ) "), (r"\!", " ! "), (r"\?", " ? "), (r"\;", " "), (r"\:", " "), (r"\s+", " ")]
_BASIC = [(re.compile(p), r) for p, r in _BASIC_PATTERNS]


def _basic_english(line: str):
    line = line.lower()
    for pat, rep in _BASIC:
        line = pat.sub(rep, line)
    return line.split()


_PUNCT = re.compile(r"\w+|[^\w\s]", re.UNICODE)


def get_tokenizer(name="basic_english", language=None):
    if name == "basic_english":
        return _basic_english
    if name is None:
        return str.split
    if callable(name):
        return name
    return lambda s: _PUNCT.findall(s)  # 'spacy' & friends


class Vocab:
    def __init__(self, itos):
        self.itos = list(itos)
        self.stoi = {t: i for i, t in enumerate(self.itos)}
        self.default_index = None

    def set_default_index(self, idx):
        self.default_index = idx

    def __len__(self):
        return len(self.itos)

    def __contains__(self, tok):
        return tok in self.stoi

    def __getitem__(self, tok):
        i = self.stoi.get(tok)
        if i is None:
            if self.default_index is None:
                raise RuntimeError(f"Token {tok} not found and default index is not set")
            return self.default_index
        return i

    def lookup_indices(self, tokens):
        return [self[t] for t in tokens]

    def lookup_tokens(self, indices):
        return [self.itos[i] for i in indices]

    def get_stoi(self):
        return dict(self.stoi)

    def get_itos(self):
        return list(self.itos)

    def __call__(self, tokens):
        return self.lookup_indices(tokens)


def build_vocab_from_iterator(iterator, min_freq: int = 1, specials=None, special_first: bool = True, max_tokens=None):
    counter = Counter()
    for tokens in iterator:
        counter.update(tokens)
    specials = list(specials or [])
    items = sorted(counter.items(), key=lambda kv: (-kv[1], kv[0]))  # frequency desc, then lexicographic (torchtext)
    if max_tokens is not None:
        items = items[:max_tokens - len(specials)]
    toks = [t for t, c in items if c >= min_freq and t not in specials]
    return Vocab(specials + toks if special_first else toks + specials)


class VocabTransform:
    def __init__(self, vocab):
        self.vocab = vocab

    def __call__(self, batch):
        if batch and isinstance(batch[0], (list, tuple)):
            return [self.vocab.lookup_indices(s) for s in batch]
        return self.vocab.lookup_indices(batch)


class AddToken:
    def __init__(self, token, begin: bool = True):
        self.token, self.begin = token, begin

    def __call__(self, batch):
        nested = batch and isinstance(batch[0], (list, tuple))
        rows = batch if nested else [batch]
        out = [([self.token] + list(r)) if self.begin else (list(r) + [self.token]) for r in rows]
        return out if nested else out[0]


class Truncate:
    def __init__(self, max_seq_len: int):
        self.max_seq_len = max_seq_len

    def __call__(self, batch):
        nested = batch and isinstance(batch[0], (list, tuple))
        rows = batch if nested else [batch]
        out = [list(r[:self.max_seq_len]) for r in rows]
        return out if nested else out[0]


class ToTensor:
    def __init__(self, padding_value=None, dtype=torch.long):
        self.padding_value, self.dtype = padding_value, dtype

    def __call__(self, batch):
        if batch and isinstance(batch[0], (list, tuple)):
            if self.padding_value is None:
                return torch.tensor(batch, dtype=self.dtype)
            m = max(len(r) for r in batch)
            return torch.tensor([list(r) + [self.padding_value] * (m - len(r)) for r in batch], dtype=self.dtype)
        return torch.tensor(batch, dtype=self.dtype)


class PadTransform:
    def __init__(self, max_length: int, pad_value: int):
        self.max_length, self.pad_value = max_length, pad_value

    def __call__(self, x: torch.Tensor):
        n = x.size(-1)
        if n < self.max_length:
            x = torch.nn.functional.pad(x, (0, self.max_length - n), value=self.pad_value)
        return x


class Sequential:
    def __init__(self, *transforms):
        self.transforms = transforms

    def __call__(self, x):
        for t in self.transforms:
            x = t(x)
        return x
