"""Loss plotter — the script form of /root/reference/plot.ipynb (cells 0-1): parses
``"{step} {train|val} {value}"`` lines and saves train/val curves to a PNG."""
import sys


def parse_log(path):
    train, val = [], []
    with open(path) as f:
        for line in f:
            parts = line.split()
            if len(parts) == 3:
                step, mode, err = int(parts[0]), parts[1], float(parts[2])
                if mode == "train":
                    train.append(err)
                elif mode == "val":
                    val.append((step, err))
    val.sort(key=lambda x: x[0])
    return train, val


def main(log="log/log.txt", out="log/validation_loss.png"):
    train, val = parse_log(log)
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except ImportError:
        print(f"matplotlib not installed; parsed {len(train)} train / {len(val)} val points")
        return train, val
    plt.figure(figsize=(8, 6))
    plt.plot(range(len(train)), train, label="Training Error")
    plt.plot([v[0] for v in val], [v[1] for v in val], label="Validation Loss", color="red", marker="x")
    plt.xlabel("Step (#)")
    plt.ylabel("Error/Loss")
    plt.title("Training and Validation Errors/Losses")
    plt.legend()
    plt.savefig(out)
    return train, val


if __name__ == "__main__":
    main(*sys.argv[1:])
