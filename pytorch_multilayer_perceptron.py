"""Single-process MLP baseline — the reference's ``pytorch_multilayer_perceptron.py`` on b200ml.

Spark is used only to read the libsvm sample (pytorch_multilayer_perceptron.py:24-30); MLP 4-5-4-3 with sigmoids,
CrossEntropyLoss, SGD(lr 0.03), 100 epochs, batch 30, 60/40 random split, then last-batch test loss/accuracy.
``B200ML_MLP_PRESET=mnist784`` switches to the BASELINE.json preset (synthetic 784-dim MNIST-shape batches).
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import torch  # noqa: E402
from torch import nn  # noqa: E402
from torch.utils.data import DataLoader, TensorDataset, random_split  # noqa: E402

from b200ml.data.libsvm import sample_multiclass_path  # noqa: E402
from b200ml.models import Multilayer_perceptor, mnist_mlp  # noqa: E402
from b200ml.ops.loss import cross_entropy  # noqa: E402
from b200ml.ops.metrics import argmax_accuracy  # noqa: E402
from b200ml.ops.optim import FlatOptimizer  # noqa: E402
from b200ml.spark import SparkSession  # noqa: E402
from b200ml.utils import script as S  # noqa: E402

# environment contract kept from the reference (pytorch_multilayer_perceptron.py:15-21)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "12355")
os.environ.setdefault("WORLD_SIZE", "1")
os.environ.setdefault("RANK", "0")
os.environ.setdefault("OMP_NUM_THREADS", "1")


def accuracy_fn(labels, predictions):
    correct = torch.eq(labels, predictions).sum().item()
    return (correct / len(predictions)) * 100


def load_data():
    if os.environ.get("B200ML_MLP_PRESET") == "mnist784":
        g = torch.Generator().manual_seed(0)
        n = 600 if S.quick() else 6000
        y = torch.randint(0, 10, (n,), generator=g)
        proto = torch.randn(10, 784, generator=g)
        return proto[y] * 0.5 + torch.randn(n, 784, generator=g), y, mnist_mlp()
    spark = SparkSession.builder.appName("MLP_TORCH").config("spark.executor.cores", "2").getOrCreate()
    train_data = spark.read.format("libsvm").load(sample_multiclass_path())
    pandas_data = train_data.toPandas()
    X = torch.stack(pandas_data["features"].apply(lambda v: torch.tensor(v.toArray(), dtype=torch.float32)).tolist())
    y = torch.tensor(data=pandas_data["label"].values, dtype=torch.long)
    spark.stop()
    return X, y, Multilayer_perceptor()


def main():
    device = S.pick_device()
    features_tensor, target_tensor, multilayer_model = load_data()
    dataset = TensorDataset(features_tensor, target_tensor)
    train_size = int(0.6 * len(dataset))
    train_dataset, test_dataset = random_split(dataset, [train_size, len(dataset) - train_size])
    train_loader = DataLoader(train_dataset, batch_size=30, shuffle=True)
    test_loader = DataLoader(test_dataset, batch_size=30, shuffle=False)

    multilayer_model = multilayer_model.to(device)
    optimizer = FlatOptimizer(multilayer_model.parameters(), kind="sgd", lr=0.03)
    epochs = S.env_int("B200ML_EPOCHS", 10 if S.quick() else 100)
    multilayer_model.train()
    timer = S.DeviceTimer(device).start()
    for epoch in range(epochs):
        for features, labels in train_loader:
            features, labels = features.to(device), labels.to(device)
            predictions = multilayer_model(features)
            loss = cross_entropy(predictions, labels)
            optimizer.zero_grad()
            loss.backward()
            optimizer.step()
    print(f"PyTorch Training Time: {timer.elapsed():.2f} seconds")

    multilayer_model.eval()
    total_test_loss, correct, count = 0.0, None, 0
    with torch.no_grad():
        for features, labels in test_loader:
            features, labels = features.to(device), labels.to(device)
            predictions = multilayer_model(features)
            predicted_classes, correct = argmax_accuracy(predictions, labels, correct)   # device-side arg-max + count
            test_loss = nn.functional.cross_entropy(predictions.float(), labels)
            total_test_loss += test_loss.item()
            count += len(labels)
    print("Test loss on last batch:", test_loss.item())
    print("accuracy is:", accuracy_fn(labels, predicted_classes))       # last batch, as the reference prints
    print("full test-set accuracy:", 100.0 * int(correct) / max(count, 1))
    return multilayer_model


if __name__ == "__main__":
    main()
