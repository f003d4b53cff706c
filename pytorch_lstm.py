"""Single-process LSTM text classifier — the reference's ``pytorch_lstm.py`` on b200ml.

AG_NEWS-shaped (label 1..4, text) pairs → basic_english tokenizer → vocab (4 specials first) → VocabTransform /
AddToken(1) / Truncate(128) / AddToken(2) / ToTensor(pad 0) → LSTM(V, 32, 32, 4, 2 layers) → CE on the last position →
Adam(1e-3), 3 epochs, batch 32; accuracy printed every 100 batches.  AG_NEWS needs a network, so the synthetic corpus is
used unless ``B200ML_AGNEWS`` points at a CSV of ``label,text`` lines.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import torch  # noqa: E402
from torch.utils.data import DataLoader  # noqa: E402

from b200ml.data import text as T  # noqa: E402
from b200ml.data.synthetic import SyntheticNews  # noqa: E402
from b200ml.models import LSTM  # noqa: E402
from b200ml.ops.loss import cross_entropy  # noqa: E402
from b200ml.ops.metrics import argmax_accuracy  # noqa: E402
from b200ml.ops.optim import FlatOptimizer  # noqa: E402
from b200ml.utils import script as S  # noqa: E402

learning_rate, epochs, batch_size, max_len, hidden_size, num_layers, output_dim = 0.001, 3, 32, 128, 32, 2, 4


def accuracy_fn(predictions, labels):
    correct = torch.eq(predictions, labels).sum().item()
    return (correct / len(predictions)) * 100


def load_dataset():
    path = os.environ.get("B200ML_AGNEWS")
    if path and os.path.isfile(path):
        with open(path, encoding="utf-8") as fh:
            return [(int(line.split(",", 1)[0].strip('"')), line.split(",", 1)[1].strip()) for line in fh if "," in line]
    return SyntheticNews(n=320 if S.quick() else 4000, vocab_words=500 if S.quick() else 5000).items


def build_pipeline(train_dataset):
    tokenizer = T.get_tokenizer("basic_english")
    vocab = T.build_vocab_from_iterator((tokenizer(text) for _, text in train_dataset), min_freq=1,
                                        specials=["<pad>", "<sos>", "<eos>", "<unk>"], special_first=True)
    vocab.set_default_index(vocab["<unk>"])
    text_tranform = T.Sequential(T.VocabTransform(vocab=vocab), T.AddToken(1, begin=True), T.Truncate(max_seq_len=max_len),
                                 T.AddToken(2, begin=False), T.ToTensor(padding_value=0))
    return tokenizer, vocab, text_tranform


def main():
    device = S.pick_device()
    train_dataset = load_dataset()
    tokenizer, vocab, text_tranform = build_pipeline(train_dataset)
    text_classifier = LSTM(vocab_size=len(vocab), embedding_dim=hidden_size, hidden_size=hidden_size,
                           output_size=output_dim, num_layers=num_layers).to(device)
    optimizer = FlatOptimizer(text_classifier.parameters(), kind="adam", lr=learning_rate)
    loader = DataLoader(train_dataset, batch_size=batch_size, shuffle=True, drop_last=True)
    n_epochs = S.env_int("B200ML_EPOCHS", 1 if S.quick() else epochs)
    text_classifier.train()
    timer = S.DeviceTimer(device).start()
    for epoch in range(n_epochs):
        for batch_idx, (labels, texts) in enumerate(loader):
            bs = labels.shape[0]
            text_tokens = text_tranform([tokenizer(t) for t in texts]).to(device)
            labels = (labels - 1).to(device)
            hidden = torch.zeros(num_layers, bs, hidden_size, device=device)
            memory = torch.zeros(num_layers, bs, hidden_size, device=device)
            pred, hidden, memory = text_classifier(text_tokens, hidden, memory)
            loss = cross_entropy(pred[:, -1, :], labels)
            optimizer.zero_grad()
            loss.backward()
            optimizer.step()
            if batch_idx % 100 == 0:
                acc = accuracy_fn(argmax_accuracy(pred[:, -1, :])[0], labels)
                print(f"epoch {epoch} batch {batch_idx} loss {loss.item():.4f} accuracy {acc:.1f}")
    print(f"Training time is : {timer.elapsed()}")
    return text_classifier


if __name__ == "__main__":
    main()
