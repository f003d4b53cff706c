"""Single-process CNN baseline — the reference's ``pytorch_cnn.py`` on b200ml.

FashionMNISTModel(1, 10, 10), batch 32, 3 epochs, SGD(lr 0.01), CrossEntropyLoss, then eval_func (total test loss,
accuracy).  FashionMNIST needs a network, so the data is the synthetic FashionMNIST-shaped set unless torchvision finds
the real files under ./data.  ``B200ML_CNN_PRESET=cifar`` switches to the BASELINE.json 3×32×32 preset.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import torch  # noqa: E402
from torch import nn  # noqa: E402
from torch.utils.data import DataLoader  # noqa: E402

from b200ml.data.synthetic import SyntheticImages  # noqa: E402
from b200ml.models import FashionMNISTModel  # noqa: E402
from b200ml.ops.loss import cross_entropy  # noqa: E402
from b200ml.ops.metrics import argmax_accuracy  # noqa: E402
from b200ml.ops.optim import FlatOptimizer  # noqa: E402
from b200ml.utils import script as S  # noqa: E402


def load_datasets():
    if os.environ.get("B200ML_CNN_PRESET") == "cifar":
        n = 256 if S.quick() else 4096
        return SyntheticImages(n, 3, 32, train=True), SyntheticImages(n // 4, 3, 32, train=False), 3, 32
    try:
        from torchvision import datasets
        from torchvision.transforms import ToTensor

        tr = datasets.FashionMNIST(root="data", train=True, download=False, transform=ToTensor())
        te = datasets.FashionMNIST(root="data", train=False, download=False, transform=ToTensor())
        return tr, te, 1, 28
    except Exception:  # noqa: BLE001  (no dataset on disk, no network)
        n = 256 if S.quick() else 4096
        return SyntheticImages(n, 1, 28, train=True), SyntheticImages(n // 4, 1, 28, train=False), 1, 28


def num_classes(dataset):
    # (kept out of train_func: cloudpickle mistakes the attribute name `classes` for the torch.classes submodule)
    return len(getattr(dataset, "classes"))


def accuracy_fn(predictions, labels):
    correct = torch.eq(predictions, labels).sum().item()
    return (correct / len(predictions)) * 100


def main():
    device = S.pick_device()
    train_data, test_data, channels, size = load_datasets()
    BATCH_SIZE = 32
    train_dataloader = DataLoader(train_data, batch_size=BATCH_SIZE, shuffle=True)
    test_dataloader = DataLoader(test_data, batch_size=BATCH_SIZE, shuffle=False)
    class_names = train_data.classes
    cnn_model = FashionMNISTModel(input_shape=channels, hidden_units=10, output_shape=len(class_names),
                                  image_size=size).to(device)
    Epochs = S.env_int("B200ML_EPOCHS", 1 if S.quick() else 3)
    optimizer = FlatOptimizer(cnn_model.parameters(), kind="sgd", lr=0.01)

    def train_func():
        timer = S.DeviceTimer(device).start()
        cnn_model.train()
        for epoch in range(Epochs):
            total_train_loss = torch.zeros((), device=device)
            for batch_idx, (data, target) in enumerate(train_dataloader):
                data, target = data.to(device), target.to(device)
                predictions = cnn_model(data)
                train_loss = cross_entropy(predictions, target)
                optimizer.zero_grad()
                train_loss.backward()
                optimizer.step()
                total_train_loss += train_loss.detach()
        print(f"Total_train_loss is:{total_train_loss.item()}")
        print(f"Training time is : {timer.elapsed()}")

    def eval_func():
        cnn_model.eval()
        total_test_loss, correct, count = 0.0, None, 0
        with torch.no_grad():
            for data, target in test_dataloader:
                data, target = data.to(device), target.to(device)
                predictions = cnn_model(data)
                predicted_classes, correct = argmax_accuracy(predictions, target, correct)   # device-side arg-max + count
                total_test_loss += nn.functional.cross_entropy(predictions.float(), target).item()
                count += len(target)
        print(f"total_test_loss is:{total_test_loss}")
        print(f"accuracy is :{accuracy_fn(predicted_classes, target)}")   # last batch, as the reference computes it
        print(f"full test-set accuracy is :{100.0 * int(correct) / max(count, 1)}")

    train_func()
    eval_func()
    return cnn_model


if __name__ == "__main__":
    main()
