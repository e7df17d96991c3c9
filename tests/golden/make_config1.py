"""Packs the reference's own examples/binary_classification fixture (BASELINE.json configs[0]) into one small .npz so
that the GPU box -- where /root/reference does not exist -- can train on the real data:
    python tests/golden/make_config1.py        (run in the container that has /root/reference)
Origin: /root/reference/examples/binary_classification/{binary.train, binary.train.weight, binary.test,
binary.test.weight, train.conf}.  Column 0 of the .train / .test files is the label, the other 28 are the features
(tab separated, parsed as float64 here); train.conf's learning parameters are kept verbatim as one string."""
import os
import numpy as np

SRC = "/root/reference/examples/binary_classification"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "config1_binary_classification.npz")


def conf_params(path):
    keep = []
    for line in open(path):
        line = line.split("#")[0].strip()
        if not line or "=" not in line:
            continue
        k, v = [t.strip() for t in line.split("=", 1)]
        # file names, the CLI task and the socket settings are not learning parameters
        if k in ("task", "data", "valid_data", "output_model", "machine_list_file", "local_listen_port", "num_machines",
                 "use_two_round_loading", "is_save_binary_file", "num_trees", "metric_freq", "is_training_metric"):
            continue
        keep.append(f"{k}={v}")
    return " ".join(keep)


def main():
    tr = np.loadtxt(os.path.join(SRC, "binary.train"), dtype=np.float64)
    te = np.loadtxt(os.path.join(SRC, "binary.test"), dtype=np.float64)
    wtr = np.loadtxt(os.path.join(SRC, "binary.train.weight"), dtype=np.float64).astype(np.float32)
    wte = np.loadtxt(os.path.join(SRC, "binary.test.weight"), dtype=np.float64).astype(np.float32)
    params = conf_params(os.path.join(SRC, "train.conf"))
    np.savez_compressed(OUT, X_train=tr[:, 1:], y_train=tr[:, 0].astype(np.float32), w_train=wtr,
                        X_test=te[:, 1:], y_test=te[:, 0].astype(np.float32), w_test=wte, params=np.array(params))
    print(OUT, tr.shape, te.shape, params)


if __name__ == "__main__":
    main()
