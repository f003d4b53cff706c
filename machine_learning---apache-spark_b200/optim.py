"""``b200ml.optim``: ``SGD`` / ``Adam`` with the ``torch.optim`` call shape the reference uses
(``SGD(model.parameters(), lr=0.01)``, ``Adam(model.parameters(), lr=1e-3)``; ``zero_grad()`` / ``step()``), driving
the fused flat optimizer - through the data-parallel engine when the parameters belong to a
``b200ml.parallel.DistributedDataParallel`` module.  See parallel/compat.py."""
from .parallel.compat import SGD, Adam  # noqa: F401
