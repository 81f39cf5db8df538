"""Feeds the streaming object-detection example: every `--interval` seconds a new text file listing a few image paths appears in
`--streaming_path` (reference: pyzoo/zoo/examples/streaming/objectdetection/image_path_writer.py, which a Spark `textFileStream`
watches).  Files are written under a temporary name and renamed, so a watcher never reads a half-written list."""
import argparse
import os
import time


def package_path_to_text(streaming_path, file_path, batch=10, delay=3.0):
    """Write the paths of the images under `file_path`, `batch` per text file, one file every `delay` seconds."""
    os.makedirs(streaming_path, exist_ok=True)
    images = sorted(os.path.join(file_path, f) for f in os.listdir(file_path) if f.lower().endswith((".jpg", ".jpeg", ".png", ".bmp")))
    written = []
    for k in range(0, len(images), batch):
        name = os.path.join(streaming_path, "paths-%05d.txt" % (k // batch))
        with open(name + ".tmp", "w") as f:
            f.write("\n".join(images[k:k + batch]) + "\n")
        os.replace(name + ".tmp", name)
        written.append(name)
        if k + batch < len(images):
            time.sleep(delay)
    return written


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--img_path", required=True, help="folder with the images to announce")
    ap.add_argument("--streaming_path", default="/tmp/zoo/streaming")
    ap.add_argument("--batch", type=int, default=10)
    ap.add_argument("--interval", type=float, default=3.0)
    a = ap.parse_args()
    print("wrote", package_path_to_text(a.streaming_path, a.img_path, a.batch, a.interval))
