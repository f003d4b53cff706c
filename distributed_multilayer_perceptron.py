"""Spark-launched data-parallel MLP — the reference's ``distributed_multilayer_perceptron.py`` on b200ml.

    python distributed_multilayer_perceptron.py --conf spark.executor.instances=2

SparkConf/SparkSession read ``spark.executor.instances`` (default: one rank per visible GPU), the libsvm sample is read
through the session, and ``TorchDistributor(num_processes, local_mode=True, use_gpu).run(train_func)`` starts one rank
per GPU.  Inside ``train_func``: process group → model → DataParallelEngine (the DDP(model) + SGD(lr 0.03) of the
reference, with the gradient all-reduce actually happening — SURVEY §9 Q1) → 100 epochs over this rank's shard
(DistributedSampler built from the real RANK/WORLD_SIZE, Q3) → every rank evaluates.  Rank 0 returns the state_dict.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import torch  # noqa: E402
from torch import nn  # noqa: E402
from torch.utils.data import DataLoader, DistributedSampler, TensorDataset, random_split  # noqa: E402

from b200ml.data.libsvm import sample_multiclass_path  # noqa: E402
from b200ml.models import Multilayer_perceptor  # noqa: E402
from b200ml.ops.loss import cross_entropy  # noqa: E402
from b200ml.ops.metrics import argmax_accuracy  # noqa: E402
from b200ml.parallel.engine import DataParallelEngine  # noqa: E402
from b200ml.spark import SparkConf, SparkSession, TorchDistributor  # noqa: E402
from b200ml.utils import script as S  # noqa: E402


def accuracy_fn(labels, predictions):
    correct = torch.eq(labels, predictions).sum().item()
    return (correct / len(predictions)) * 100


def train_func(features_tensor, target_tensor):
    device, rank, world = S.init_worker()
    torch.manual_seed(0)  # same split on every rank
    dataset = TensorDataset(features_tensor, target_tensor)
    train_size = int(0.6 * len(dataset))
    train_dataset, test_dataset = random_split(dataset, [train_size, len(dataset) - train_size])
    sampler = DistributedSampler(train_dataset, num_replicas=world, rank=rank)
    train_loader = DataLoader(train_dataset, batch_size=30, sampler=sampler)
    test_loader = DataLoader(test_dataset, batch_size=30, shuffle=False)

    multilayer_model = Multilayer_perceptor().to(device)
    DDP_model = DataParallelEngine(multilayer_model, optimizer="sgd", lr=0.03)
    epochs = S.env_int("B200ML_EPOCHS", 10 if S.quick() else 100)
    multilayer_model.train()

    def one_step(features, labels):
        predictions = multilayer_model(features)
        loss = cross_entropy(predictions, labels)
        DDP_model.zero_grad()
        loss.backward()
        DDP_model.step()
        return loss.detach()

    train_step = DDP_model.graphed(one_step)
    timer = S.DeviceTimer(device).start()
    for epoch in range(epochs):
        sampler.set_epoch(epoch)
        for features, labels in train_loader:
            features, labels = features.to(device), labels.to(device)
            loss = train_step(features, labels)
    secs = S.max_over_ranks(timer.elapsed(), device)
    if rank == 0:
        print(f"PyTorch Training Time: {secs:.2f} seconds")
    DDP_model.weights_hash()  # every rank must hold identical weights
    DDP_model.close()
    S.finish_worker()

    multilayer_model.eval()
    total_test_loss, correct = 0, None
    with torch.no_grad():
        for features, labels in test_loader:
            features, labels = features.to(device), labels.to(device)
            predictions = multilayer_model(features)
            predicted_classes, correct = argmax_accuracy(predictions, labels, correct)   # device-side arg-max + count
            test_loss = nn.functional.cross_entropy(predictions.float(), labels)
            total_test_loss += test_loss.item()
    print("Test loss on last batch:", test_loss.item())
    print("accuracy is:", accuracy_fn(labels, predicted_classes))
    return {k: v.detach().cpu() for k, v in multilayer_model.state_dict().items()} if rank == 0 else None


if __name__ == "__main__":
    empty_config = SparkConf()
    spark = SparkSession.builder.config(conf=empty_config).getOrCreate()
    executors_n = int(spark.sparkContext.getConf().get("spark.executor.instances"))
    train_data = spark.read.format("libsvm").load(sample_multiclass_path())
    pandas_data = train_data.toPandas()
    X_train = torch.stack(pandas_data["features"].apply(lambda v: torch.tensor(v.toArray(), dtype=torch.float32)).tolist())
    target_tensor = torch.tensor(data=pandas_data["label"].values, dtype=torch.long)

    distributor = TorchDistributor(num_processes=executors_n, local_mode=True, use_gpu=torch.cuda.is_available())
    model = distributor.run(train_func, X_train, target_tensor)
    print("returned state_dict keys:", list(model.keys()) if model else None)
    spark.stop()
