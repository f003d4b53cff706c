"""MLlib-style MultilayerPerceptronClassifier — the reference's ``mllib_multilayer_perceptron_classifier.py``.

Same calls: SparkSession.builder…config(executor.instances 3)…getOrCreate() → libsvm load → randomSplit([0.6, 0.4],
1234) → MultilayerPerceptronClassifier(maxIter=100, layers=[4,5,4,3], blockSize=30, solver='l-bfgs', stepSize=0.03)
.fit(train) → transform(test) → MulticlassClassificationEvaluator(metricName="accuracy").  ``fit`` shards the rows
over ``spark.executor.instances`` ranks (one per GPU when that many GPUs are visible, CPU/gloo ranks otherwise) and
replaces Spark's per-evaluation broadcast + treeAggregate by one fused GPU all-reduce (b200ml/spark/ml.py).
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from b200ml.data.libsvm import sample_multiclass_path  # noqa: E402
from b200ml.spark import (MulticlassClassificationEvaluator, MultilayerPerceptronClassifier,  # noqa: E402
                          SparkSession)

if __name__ == "__main__":
    spark = SparkSession.builder \
        .appName("spark_mlib_multilayer_perceptron_classifier") \
        .config("spark.executor.cores", "1") \
        .config("spark.executor.memory", "8g") \
        .config("spark.executor.instances", os.environ.get("B200ML_EXECUTORS", "3")) \
        .config("spark.driver.cores", "4") \
        .config("spark.driver.memory", "8g") \
        .getOrCreate()

    train_data = spark.read.format("libsvm").load(sample_multiclass_path())
    splits = train_data.randomSplit([0.6, 0.4], 1234)
    train, test = splits[0], splits[1]
    layers = [4, 5, 4, 3]
    trainer = MultilayerPerceptronClassifier(maxIter=100, layers=layers, blockSize=30, solver="l-bfgs", stepSize=0.03)
    start_time = time.time()
    model = trainer.fit(train)
    end_time = time.time()
    print(f"PyTorch Training Time: {end_time - start_time:.2f} seconds")   # label kept from the reference (Q18)
    print(f"L-BFGS iterations: {model.summary.totalIterations}, function evaluations: "
          f"{model.summary.functionEvaluations}, workers: {model.summary.numWorkers}, final objective: "
          f"{model.summary.objectiveHistory[-1]:.6f}")

    result = model.transform(test)
    predictionAndLabels = result.select("prediction", "label")
    evaluator = MulticlassClassificationEvaluator(metricName="accuracy")
    print("Test set accuracy = " + str(evaluator.evaluate(predictionAndLabels)))
    spark.stop()
