"""Synthetic datasets with the shapes of the reference's datasets (BASELINE.json: "All runs use synthetic data of the
named shape"; the real FashionMNIST / AG_NEWS / Multi30k downloads need a network this box does not have).

  * ``SyntheticImages``   FashionMNIST-shape (1×28×28, 10 classes; distributed_cnn.py:90-106) or CIFAR-shape (3×32×32)
  * ``SyntheticNews``     AG_NEWS-shape (label 1..4, text) pairs over a Zipfian synthetic vocabulary (distributed_lstm.py:69-70)
  * ``SyntheticParallelCorpus``  Multi30k-shape (en, de) sentence pairs (pytorch_machine_translator.py:14-16)
Images carry a class-dependent pattern and texts class-dependent keywords, so loss curves actually go down.
"""
from __future__ import annotations

import torch
from torch.utils.data import Dataset


class SyntheticImages(Dataset):
    classes = ["T-shirt/top", "Trouser", "Pullover", "Dress", "Coat", "Sandal", "Shirt", "Sneaker", "Bag", "Ankle boot"]

    def __init__(self, n: int = 2048, channels: int = 1, size: int = 28, num_classes: int = 10, seed: int = 0,
                 train: bool = True):
        g = torch.Generator().manual_seed(seed + (0 if train else 1))
        self.targets = torch.randint(0, num_classes, (n,), generator=g)
        proto = torch.randn(num_classes, channels, size, size, generator=torch.Generator().manual_seed(seed + 77))
        noise = torch.randn(n, channels, size, size, generator=g)
        self.data = (0.5 + 0.25 * proto[self.targets] + 0.25 * noise).clamp_(0, 1)
        self.classes = self.classes[:num_classes] if num_classes <= 10 else [str(i) for i in range(num_classes)]

    def __len__(self):
        return self.data.shape[0]

    def __getitem__(self, i):
        return self.data[i], int(self.targets[i])


def _zipf_words(n_words: int):
    return [f"w{i}" for i in range(n_words)]


class SyntheticNews(Dataset):
    """Items are (label in 1..4, text) like torchtext's AG_NEWS."""

    def __init__(self, n: int = 4000, vocab_words: int = 5000, min_len: int = 12, max_len: int = 90, seed: int = 0):
        g = torch.Generator().manual_seed(seed)
        words = _zipf_words(vocab_words)
        ranks = torch.arange(1, vocab_words + 1, dtype=torch.float64)
        probs = (1.0 / ranks) / (1.0 / ranks).sum()
        self.items = []
        lens = torch.randint(min_len, max_len + 1, (n,), generator=g)
        labels = torch.randint(1, 5, (n,), generator=g)
        for i in range(n):
            idx = torch.multinomial(probs, int(lens[i]), replacement=True, generator=g)
            toks = [words[j] for j in idx.tolist()]
            lab = int(labels[i])
            for k in range(0, len(toks), 5):           # class keywords make the task learnable
                toks[k] = f"topic{lab}k{(k // 5) % 7}"
            self.items.append((lab, " ".join(toks) + " ."))

    def __len__(self):
        return len(self.items)

    def __getitem__(self, i):
        return self.items[i]

    def __iter__(self):
        return iter(self.items)


class SyntheticParallelCorpus(Dataset):
    """(english, german) sentence pairs; the 'german' side is a deterministic word-for-word recoding of the english
    side plus a suffix, so a translator can fit it."""

    def __init__(self, n: int = 2900, vocab_words: int = 3000, min_len: int = 5, max_len: int = 30, seed: int = 0):
        g = torch.Generator().manual_seed(seed)
        en_words = [f"en{i}" for i in range(vocab_words)]
        ranks = torch.arange(1, vocab_words + 1, dtype=torch.float64)
        probs = (1.0 / ranks) / (1.0 / ranks).sum()
        lens = torch.randint(min_len, max_len + 1, (n,), generator=g)
        self.items = []
        for i in range(n):
            idx = torch.multinomial(probs, int(lens[i]), replacement=True, generator=g).tolist()
            en = " ".join(en_words[j] for j in idx) + " ."
            de = " ".join(f"de{(j * 7 + 3) % vocab_words}" for j in idx) + " ."
            self.items.append((en, de))

    def __len__(self):
        return len(self.items)

    def __getitem__(self, i):
        return self.items[i]

    def __iter__(self):
        return iter(self.items)


def synthetic_token_batch(batch: int, seq: int, vocab: int, generator=None, pad_tail: bool = True):
    """[batch, seq] int64 ids in [4, vocab) with <sos>=1 first, <eos>=2 last and <pad>=0 tails — the shape the
    reference's transform pipelines emit (pytorch_machine_translator.py:70-98)."""
    g = generator
    ids = torch.randint(4, vocab, (batch, seq), generator=g)
    if pad_tail:
        lens = torch.randint(max(3, seq // 8), seq + 1, (batch,), generator=g)
        ar = torch.arange(seq)[None, :]
        ids = torch.where(ar < lens[:, None], ids, torch.zeros_like(ids))
        ids[torch.arange(batch), lens - 1] = 2
    ids[:, 0] = 1
    return ids
