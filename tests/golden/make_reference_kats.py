"""Writes tests/golden/reference_kats.json: every known answer the REFERENCE'S OWN tests hold for
the K-means DAL path.  The expected values below are transcribed from the reference's assertions,
not produced by our oracle (run here, where /root/reference is mounted; the JSON travels):

  kat_fit_summary   mllib-dal/src/test/scala/org/apache/spark/ml/clustering/IntelKMeansSuite.scala:110-141
                    data: KMeansSuite.generateKMeansData(spark, 50, 3, 5) (:484-489, row i = [i % 5] * 3,
                    i = 1..50), k = 5  =>  5 centres, labels == {0..4}, trainingCost < 0.1, numIter == 1,
                    cluster sizes sum to N.  (numIter == 1 requires initial centres that already are
                    the five distinct points, which k-means|| returns on this data.)
  kat_four_centers  IntelKMeansSuite.scala:413-480: 8 points (:416-423), k = 4, maxIter = 10  =>  the
                    pairs listed at :435-442 are co-clustered, 4 distinct predictions; DAL centres ==
                    vanilla-Spark centres exactly (:479), i.e. the four pair means.
  kat_array_input   IntelKMeansSuite.scala:204-218: trainingCost identical (abs 1e-6) for Double and
                    Float input on the 50x3 fixture, k = 5, maxIter = 1  =>  FP32 storage is inside
                    the reference's own tolerance.
  kat_defaults      IntelKMeansSuite.scala:50-70: k=2, maxIter=20, initMode k-means||, initSteps=2,
                    tol=1e-4, euclidean.
  kat_sample_data   examples/data/sample_kmeans_data.txt:1-6 with examples/kmeans/src/main/scala/
                    org/apache/spark/examples/ml/KMeansExample.scala:47-65 (k=2): the example only
                    prints; the separable data has the unique fixed point centres (0.1,..),(9.1,..).
"""
import json
import os
import re

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    suite = open(os.path.join(REF, "mllib-dal/src/test/scala/org/apache/spark/ml/clustering/IntelKMeansSuite.scala")).read()
    # the 8 points of "Four centers with weightCol" (first df1 block)
    blk = suite[suite.index('test("Four centers with weightCol")'):]
    blk = blk[:blk.index("map(v => TestRow(v))")]
    pts = [[float(a), float(b)] for a, b in re.findall(r"Vectors\.dense\((-?[\d.]+), (-?[\d.]+)\)", blk)]
    assert len(pts) == 8
    sample = []
    for line in open(os.path.join(REF, "examples/data/sample_kmeans_data.txt")):
        parts = line.split()
        sample.append([float(p.split(":")[1]) for p in parts[1:]])
    assert len(sample) == 6
    assert "generateKMeansData(spark, 50, 3, 5)" in suite or "generateKMeansData" in suite
    fixture = [[float(i % 5)] * 3 for i in range(1, 51)]
    kats = {
        "kat_fit_summary": {"data": fixture, "k": 5, "expect": {"n_centers": 5, "labels_set": [0, 1, 2, 3, 4],
                                                               "cost_lt": 0.1, "num_iter": 1, "sizes_sum": 50}},
        "kat_four_centers": {"data": pts, "k": 4, "max_iter": 10,
                             "expect": {"pairs": [[0, 1], [2, 3], [4, 5], [6, 7]],
                                        "centers": [[(pts[a][0] + pts[b][0]) / 2, (pts[a][1] + pts[b][1]) / 2]
                                                    for a, b in [(0, 1), (2, 3), (4, 5), (6, 7)]]}},
        "kat_array_input": {"data": fixture, "k": 5, "max_iter": 1, "expect": {"cost_abs_tol": 1e-6}},
        "kat_defaults": {"k": 2, "maxIter": 20, "initMode": "k-means||", "initSteps": 2, "tol": 1e-4,
                         "distanceMeasure": "euclidean"},
        "kat_sample_data": {"data": sample, "k": 2, "expect": {"centers": [[0.1, 0.1, 0.1], [9.1, 9.1, 9.1]],
                                                               "labels": [0, 0, 0, 1, 1, 1], "cost": 0.12}},
    }
    with open(os.path.join(HERE, "reference_kats.json"), "w") as f:
        json.dump(kats, f)
    print("wrote", len(kats), "KATs")


if __name__ == "__main__":
    main()
