"""Spark-launched data-parallel LSTM text classifier — the reference's ``distributed_lstm.py`` on b200ml.

    python distributed_lstm.py --conf spark.executor.instances=8

Driver as in the reference (SparkConf → SparkSession → executor count → TorchDistributor.run(train_func)); the worker
builds the text pipeline, shards the corpus by rank (the reference computes a sharding_filter and then ignores it —
SURVEY §9 Q4), and trains LSTM(V, 32, 32, 4) with Adam(1e-3) through the DataParallelEngine: persistent LSTM kernels
for the recurrence and the fused all-reduce + Adam for the 12 MB embedding-dominated gradient.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import torch  # noqa: E402
from torch.utils.data import DataLoader, DistributedSampler  # noqa: E402

from b200ml.models import LSTM  # noqa: E402
from b200ml.ops.loss import cross_entropy  # noqa: E402
from b200ml.ops.metrics import argmax_accuracy  # noqa: E402
from b200ml.parallel.engine import DataParallelEngine  # noqa: E402
from b200ml.spark import SparkConf, SparkSession, TorchDistributor  # noqa: E402
from b200ml.utils import script as S  # noqa: E402
from pytorch_lstm import (batch_size, build_pipeline, epochs, hidden_size, learning_rate, load_dataset,  # noqa: E402
                          num_layers, output_dim)


def train_func():
    device, rank, world = S.init_worker()
    train_dataset = load_dataset()
    tokenizer, vocab, text_tranform = build_pipeline(train_dataset)
    torch.manual_seed(0)
    text_classifier = LSTM(vocab_size=len(vocab), embedding_dim=hidden_size, hidden_size=hidden_size,
                           output_size=output_dim, num_layers=num_layers).to(device)
    DDP_model = DataParallelEngine(text_classifier, optimizer="adam", lr=learning_rate)
    sampler = DistributedSampler(train_dataset, num_replicas=world, rank=rank, shuffle=True)
    loader = DataLoader(train_dataset, batch_size=batch_size, sampler=sampler, drop_last=True)
    n_epochs = S.env_int("B200ML_EPOCHS", 1 if S.quick() else epochs)
    text_classifier.train()

    def one_step(text_tokens, labels):
        bs = labels.shape[0]
        hidden = torch.zeros(num_layers, bs, hidden_size, device=device)
        memory = torch.zeros(num_layers, bs, hidden_size, device=device)
        pred, hidden, memory = text_classifier(text_tokens, hidden, memory)
        loss = cross_entropy(pred[:, -1, :], labels)
        DDP_model.zero_grad()
        loss.backward()
        DDP_model.step()
        return argmax_accuracy(pred[:, -1, :], labels)[1]   # device-side arg-max + number of correct rows

    # batches are padded to their own longest text (<= 129 tokens); most hit the truncation cap, so a handful of
    # sequence lengths covers nearly every batch: each gets its own CUDA graph, rare lengths run eagerly
    train_step = DDP_model.graphed(one_step, max_graphs=16)
    timer = S.DeviceTimer(device).start()
    n_samples = 0
    train_acc = torch.zeros(1, dtype=torch.int32, device=device)
    for epoch in range(n_epochs):
        sampler.set_epoch(epoch)
        for batch_idx, (labels, texts) in enumerate(loader):
            bs = labels.shape[0]
            text_tokens = text_tranform([tokenizer(t) for t in texts]).to(device)
            labels = (labels - 1).to(device)
            train_acc += train_step(text_tokens, labels)
            n_samples += bs
    secs = S.max_over_ranks(timer.elapsed(), device)
    DDP_model.weights_hash()
    if rank == 0:
        print(f"Training time is : {secs}")
        print(f"train accuracy: {100.0 * float(train_acc) / max(n_samples, 1):.2f}")
        S.log_json(rank, model="lstm", world=world, samples_per_s=n_samples * world / secs)
    DDP_model.close()
    S.finish_worker()
    return {k: v.detach().cpu() for k, v in text_classifier.state_dict().items()} if rank == 0 else None


if __name__ == "__main__":
    empty_config = SparkConf()
    spark = SparkSession.builder.config(conf=empty_config).getOrCreate()
    executors_n = int(spark.sparkContext.getConf().get("spark.executor.instances"))
    distributor = TorchDistributor(num_processes=executors_n, local_mode=False, use_gpu=torch.cuda.is_available())
    model = distributor.run(train_func)
    print("returned state_dict keys:", list(model.keys())[:4] if model else None)
    spark.stop()
