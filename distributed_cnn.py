"""Spark-launched data-parallel CNN — the reference's ``distributed_cnn.py`` on b200ml.

    python distributed_cnn.py --conf spark.executor.instances=8        (one rank per B200)

Driver: SparkConf → SparkSession → ``spark.executor.instances`` → TorchDistributor(num_processes, use_gpu).run(train_func)
→ spark.stop().  Worker: process group → FashionMNISTModel(1,10,10) → DataParallelEngine (DDP + SGD lr 0.01 with the
fused all-reduce) → 3 epochs of batch 32 over this rank's shard of the TRAIN set (the reference iterates the test
loader and pins every rank to shard 0 of 2 — SURVEY §9 Q2/Q3) → rank 0 evaluates and returns the state_dict.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import torch  # noqa: E402
from torch import nn  # noqa: E402
from torch.utils.data import DataLoader, DistributedSampler  # noqa: E402

from b200ml.models import FashionMNISTModel  # noqa: E402
from b200ml.ops.loss import cross_entropy  # noqa: E402
from b200ml.ops.metrics import argmax_accuracy  # noqa: E402
from b200ml.parallel.engine import DataParallelEngine  # noqa: E402
from b200ml.spark import SparkConf, SparkSession, TorchDistributor  # noqa: E402
from b200ml.utils import script as S  # noqa: E402
from pytorch_cnn import accuracy_fn, load_datasets, num_classes  # noqa: E402

BATCH_SIZE = 32


def train_func():
    device, rank, world = S.init_worker()
    train_data, test_data, channels, size = load_datasets()
    sampler = DistributedSampler(train_data, num_replicas=world, rank=rank)
    train_dataloader = DataLoader(train_data, batch_size=BATCH_SIZE, shuffle=False, sampler=sampler, drop_last=True)
    test_dataloader = DataLoader(test_data, batch_size=BATCH_SIZE, shuffle=False)
    torch.manual_seed(0)
    cnn_model = FashionMNISTModel(input_shape=channels, hidden_units=10, output_shape=num_classes(train_data),
                                  image_size=size).to(device)
    DDP_model = DataParallelEngine(cnn_model, optimizer="sgd", lr=0.01)
    Epochs = S.env_int("B200ML_EPOCHS", 1 if S.quick() else 3)
    def one_step(data, target):
        predictions = cnn_model(data)
        train_loss = cross_entropy(predictions, target)
        DDP_model.zero_grad()
        train_loss.backward()
        DDP_model.step()
        return train_loss.detach()

    train_step = DDP_model.graphed(one_step)   # full batches replay as one CUDA graph; a short last batch runs eagerly
    timer = S.DeviceTimer(device).start()
    cnn_model.train()
    n_samples = 0
    for epoch in range(Epochs):
        sampler.set_epoch(epoch)
        total_train_loss = torch.zeros((), device=device)
        for batch_idx, (data, target) in enumerate(train_dataloader):
            data, target = data.to(device, non_blocking=True), target.to(device, non_blocking=True)
            total_train_loss += train_step(data, target)
            n_samples += len(target)
    secs = S.max_over_ranks(timer.elapsed(), device)
    DDP_model.weights_hash()
    if rank == 0:
        print(f"Total_train_loss is:{total_train_loss.item()}")
        print(f"Training time is : {secs}")
        S.log_json(rank, model="cnn", world=world, samples_per_s=n_samples * world / secs)
        eval_func(cnn_model, test_dataloader, device)
    DDP_model.close()
    S.finish_worker()
    return {k: v.detach().cpu() for k, v in cnn_model.state_dict().items()} if rank == 0 else None


def eval_func(model, test_dataloader, device):
    model.eval()
    total_test_loss, correct, count = 0.0, None, 0
    with torch.no_grad():
        for data, target in test_dataloader:
            data, target = data.to(device), target.to(device)
            predictions = model(data)
            predicted_classes, correct = argmax_accuracy(predictions, target, correct)   # device-side arg-max + count
            total_test_loss += nn.functional.cross_entropy(predictions.float(), target).item()
            count += len(target)
    print(f"total_test_loss is:{total_test_loss}")
    print(f"accuracy is :{accuracy_fn(predicted_classes, target)}")
    print(f"full test-set accuracy is :{100.0 * int(correct) / max(count, 1)}")


if __name__ == "__main__":
    empty_config = SparkConf()
    spark = SparkSession.builder.config(conf=empty_config).getOrCreate()
    executors_n = int(spark.sparkContext.getConf().get("spark.executor.instances"))
    distributor = TorchDistributor(num_processes=executors_n, local_mode=False, use_gpu=torch.cuda.is_available())
    model = distributor.run(train_func)
    print("returned state_dict keys:", list(model.keys()) if model else None)
    spark.stop()
