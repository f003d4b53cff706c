"""Spark shim, launcher env contract, data layer, MLPC vs scipy."""
import os
import sys

import numpy as np
import pytest
import torch


def test_spark_conf_sources(monkeypatch):
    from b200ml.spark import SparkConf, SparkSession
    monkeypatch.setattr(sys, "argv", ["x.py", "--conf", "spark.executor.instances=4", "--conf=spark.foo=bar"])
    monkeypatch.setenv("SPARK_CONF_spark_executor_memory", "2g")
    c = SparkConf()
    assert c.get("spark.executor.instances") == "4" and c.get("spark.foo") == "bar"
    assert c.get("spark.executor.memory") == "2g" and c.get("missing", "d") == "d"
    spark = SparkSession.builder.appName("t").config("spark.driver.cores", "4").config(conf=c).getOrCreate()
    assert int(spark.sparkContext.getConf().get("spark.executor.instances")) == 4
    assert spark.sparkContext.getConf().get("spark.driver.cores") == "4"
    spark.stop()
    monkeypatch.setattr(sys, "argv", ["x.py"])
    assert int(SparkConf().get("spark.executor.instances")) >= 1      # defaults to visible GPUs, else 1


def _probe(tag):
    import os as _os
    keys = ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "OMP_NUM_THREADS")
    out = {k: _os.environ.get(k) for k in keys}
    out["tag"] = tag
    print("probe", out["RANK"])
    return out


def _boom():
    import os as _os
    if _os.environ["RANK"] == "1":
        raise ValueError("rank one exploded")
    import time
    time.sleep(30)


def test_torch_distributor_env_contract_and_return(capfd):
    from b200ml.spark import TorchDistributor
    out = TorchDistributor(num_processes=2, local_mode=True, use_gpu=False).run(_probe, "hello")
    assert out["RANK"] == "0" and out["WORLD_SIZE"] == "2" and out["MASTER_ADDR"] == "127.0.0.1"
    assert out["tag"] == "hello" and int(out["MASTER_PORT"]) > 0 and out["OMP_NUM_THREADS"]
    txt = capfd.readouterr().out
    assert "[rank 0] probe 0" in txt and "[rank 1] probe 1" in txt


def test_torch_distributor_fails_fast():
    import time
    from b200ml.spark import TorchDistributor
    t0 = time.time()
    with pytest.raises(RuntimeError, match="(?s)rank 1 exited.*rank one exploded"):
        TorchDistributor(num_processes=2, local_mode=True, use_gpu=False).run(_boom)
    assert time.time() - t0 < 25      # the sleeping rank was torn down, not waited for


def test_libsvm_dataframe_pipeline(tmp_path, monkeypatch):
    monkeypatch.setenv("B200ML_DATA", str(tmp_path))
    monkeypatch.delenv("SPARK_HOME", raising=False)
    from b200ml.data.libsvm import load_libsvm, sample_multiclass_path, save_libsvm
    from b200ml.spark import SparkSession
    path = sample_multiclass_path()
    x, y = load_libsvm(path)
    assert x.shape == (150, 4) and set(y.tolist()) == {0.0, 1.0, 2.0} and np.abs(x).max() <= 1.0
    save_libsvm(str(tmp_path / "t.txt"), np.array([[0, 1.5, 0], [2, 0, 0]]), np.array([1, 0]))
    x2, y2 = load_libsvm(str(tmp_path / "t.txt"), num_features=3)
    assert x2.tolist() == [[0, 1.5, 0], [2, 0, 0]] and y2.tolist() == [1, 0]
    df = SparkSession.builder.getOrCreate().read.format("libsvm").load(path)
    a, b = df.randomSplit([0.6, 0.4], 1234)
    assert a.count() + b.count() == 150 and 60 < a.count() < 120
    a2, _ = df.randomSplit([0.6, 0.4], 1234)
    assert a2.count() == a.count()                                      # seeded
    pdf = df.toPandas()
    assert list(pdf.columns) == ["label", "features"] and pdf["features"][0].toArray().shape == (4,)
    assert df.select("label").columns == ["label"]


def test_text_pipeline_shapes():
    from b200ml.data import text as T
    from b200ml.data.synthetic import SyntheticNews, SyntheticParallelCorpus, synthetic_token_batch
    tok = T.get_tokenizer("basic_english")
    assert tok("Hello, World! It's (fine).") == ["hello", ",", "world", "!", "it", "'", "s", "(", "fine", ")", "."]
    ds = SyntheticNews(40, vocab_words=200)
    v = T.build_vocab_from_iterator((tok(t) for _, t in ds), min_freq=1, specials=["<pad>", "<sos>", "<eos>", "<unk>"])
    v.set_default_index(v["<unk>"])
    assert [v[s] for s in ("<pad>", "<sos>", "<eos>", "<unk>")] == [0, 1, 2, 3] and v["never-seen"] == 3
    tr = T.Sequential(T.VocabTransform(v), T.AddToken(1, begin=True), T.Truncate(128), T.AddToken(2, begin=False),
                      T.ToTensor(padding_value=0))
    out = tr([tok(ds[0][1]), tok(ds[1][1]), ["w1"] * 300])
    assert out.shape == (3, 129) and out[2, 0] == 1 and out[2, -1] == 2          # length <= 129
    pad = T.PadTransform(200, 0)(out)
    assert pad.shape == (3, 200)
    en, de = SyntheticParallelCorpus(10)[0]
    assert len(en.split()) == len(de.split())
    b = synthetic_token_batch(4, 50, 100, torch.Generator().manual_seed(0))
    assert b.shape == (4, 50) and (b[:, 0] == 1).all()


def test_mlpc_layout_objective_and_scipy_agreement(tmp_path, monkeypatch):
    scipy_opt = pytest.importorskip("scipy.optimize")
    monkeypatch.setenv("B200ML_DATA", str(tmp_path))
    from b200ml.data.libsvm import load_libsvm, sample_multiclass_path
    from b200ml.spark import ml
    X, y = load_libsvm(sample_multiclass_path())
    layers = [4, 5, 4, 3]
    assert ml.num_weights(layers) == 64
    w0 = ml.init_weights(layers, 7)
    assert w0.dtype == torch.float64 and w0[:25].abs().max() <= 2.4 / 2.0 + 1e-9      # U(-2.4,2.4)/sqrt(4)
    Xt, yt = torch.tensor(X), torch.tensor(y, dtype=torch.long)
    rw = torch.full((150,), 1 / 30.0, dtype=torch.float64)

    def fn(w):
        l, g = ml.loss_and_grad(torch.as_tensor(w), layers, Xt, yt, rw)
        return float(l) / 5, (g / 5).numpy()

    # analytic gradient vs autograd through the same forward
    wv = w0.clone().requires_grad_()
    _, P, _ = ml.forward(wv, layers, Xt)
    (-(rw * torch.log(P.gather(1, yt[:, None]).squeeze(1))).sum() / 5).backward()
    assert np.allclose(fn(w0.numpy())[1], wv.grad.numpy(), atol=1e-10)
    res = scipy_opt.minimize(fn, w0.numpy(), jac=True, method="L-BFGS-B", options={"maxiter": 100})
    opt = ml.LBFGS(lambda w: tuple(torch.as_tensor(v) for v in fn(w.numpy())), max_iter=100)
    w, iters = opt.minimize(w0)
    assert all(b <= a + 1e-12 for a, b in zip(opt.history, opt.history[1:]))       # objective non-increasing
    assert opt.history[-1] <= max(2 * res.fun, 1e-3) and iters <= 100


def test_mlpc_fit_transform_evaluate_single_and_two_workers(tmp_path, monkeypatch):
    monkeypatch.setenv("B200ML_DATA", str(tmp_path))
    from b200ml.data.libsvm import sample_multiclass_path
    from b200ml.spark import (MulticlassClassificationEvaluator, MultilayerPerceptronClassificationModel,
                              MultilayerPerceptronClassifier, SparkSession)
    spark = SparkSession.builder.config("spark.executor.instances", "1").getOrCreate()
    df = spark.read.format("libsvm").load(sample_multiclass_path())
    train, test = df.randomSplit([0.6, 0.4], 1234)
    kw = dict(maxIter=100, layers=[4, 5, 4, 3], blockSize=30, solver="l-bfgs", stepSize=0.03, seed=5, device="cpu")
    m1 = MultilayerPerceptronClassifier(**kw).fit(train)
    res = m1.transform(test)
    assert {"rawPrediction", "probability", "prediction"} <= set(res.columns)
    acc = MulticlassClassificationEvaluator(metricName="accuracy").evaluate(res.select("prediction", "label"))
    assert acc >= 0.9 and len(m1.weights) == 64 and m1.summary.totalIterations <= 100
    assert 0.9 <= MulticlassClassificationEvaluator().evaluate(res) <= 1.0           # default metric f1
    # equal shards of whole blocks => 2 workers compute the same objective => same iterates
    sub = train._take(np.arange(60))
    a = MultilayerPerceptronClassifier(**kw).fit(sub)
    b = MultilayerPerceptronClassifier(numWorkers=2, **kw).fit(sub)
    assert b.summary.numWorkers == 2
    assert np.allclose(a.summary.objectiveHistory[:5], b.summary.objectiveHistory[:5], rtol=1e-9)
    assert np.allclose(a.weights.toArray(), b.weights.toArray(), atol=1e-6)
    m1.save(str(tmp_path / "mlpc.pkl"))
    m2 = MultilayerPerceptronClassificationModel.load(str(tmp_path / "mlpc.pkl"))
    assert m2.predict(test.collect()[0]["features"]) == m1.predict(test.collect()[0]["features"])
    gd = MultilayerPerceptronClassifier(maxIter=50, layers=[4, 5, 4, 3], solver="gd", stepSize=0.5, seed=5, device="cpu").fit(train)
    assert gd.summary.objectiveHistory[-1] < gd.summary.objectiveHistory[0]
    spark.stop()


def test_pyspark_shim_import_paths():
    from b200ml.spark import install_pyspark_shim
    if install_pyspark_shim():
        from pyspark import SparkConf  # noqa: F401
        from pyspark.ml.classification import MultilayerPerceptronClassifier  # noqa: F401
        from pyspark.ml.evaluation import MulticlassClassificationEvaluator  # noqa: F401
        from pyspark.ml.torch.distributor import TorchDistributor  # noqa: F401
        from pyspark.sql import SparkSession  # noqa: F401


def test_argmax_accuracy_cpu_restatement():
    """ops/metrics.py on CPU tensors (the oracle of the K14 kernel test): predictions and the accumulating counter."""
    import torch

    from b200ml.ops.metrics import argmax_accuracy
    z = torch.tensor([[0.1, 0.9, 0.0], [2.0, 1.0, 2.0], [0.0, 0.0, 3.0]])
    y = torch.tensor([1, 0, 1])
    pred, correct = argmax_accuracy(z, y)
    assert pred.tolist() == [1, 0, 2] and int(correct) == 2
    _, correct = argmax_accuracy(z, y, correct)
    assert int(correct) == 4


def test_mlpc_reproducible_mode_is_bit_identical_across_worker_counts(tmp_path, monkeypatch):
    """``reproducible=True`` (SURVEY W13 "1-vs-N identical iterates"): canonical row chunks, per-chunk evaluation, fixed
    pairwise tree - the objective history and the weights must be EXACTLY the same for 1, 2 and 4 workers."""
    from b200ml.spark.dataframe import DataFrame
    from b200ml.spark.ml import MultilayerPerceptronClassifier

    rng = np.random.default_rng(7)
    n, d, k = 203, 6, 3                      # not a multiple of anything on purpose
    X = rng.normal(size=(n, d))
    y = ((X[:, 0] + 0.5 * X[:, 1] > 0).astype(int) + (X[:, 2] > 1.0)).astype(np.float64)
    df = DataFrame({"label": y}, {"features": X})
    out = {}
    for workers in (1, 2, 4):
        clf = MultilayerPerceptronClassifier(maxIter=12, layers=[d, 5, k], blockSize=16, seed=3, numWorkers=workers,
                                             device="cpu", reproducible=True)
        m = clf.fit(df)
        out[workers] = (list(m.summary.objectiveHistory), np.asarray(m.weights.toArray()))
        assert m.summary.numWorkers == workers
    for workers in (2, 4):
        assert out[workers][0] == out[1][0], (workers, out[workers][0][-3:], out[1][0][-3:])
        assert np.array_equal(out[workers][1], out[1][1])
    assert out[1][0][-1] < out[1][0][0]
    with pytest.raises(ValueError):
        MultilayerPerceptronClassifier(maxIter=2, layers=[d, 5, k], numWorkers=3, device="cpu", reproducible=True).fit(df)
