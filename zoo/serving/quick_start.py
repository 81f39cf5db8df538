"""`python -m zoo.serving.quick_start`: enqueue one image and print its result (reference: pyzoo/zoo/serving/quick_start.py)."""
import time

import numpy as np

from .client import InputQueue, OutputQueue


def run(config="config.yaml", path=None):
    inq, outq = InputQueue(config), OutputQueue(config)
    import cv2

    img = cv2.imread(path) if path else np.random.randint(0, 256, (224, 224, 3), dtype=np.uint8)
    inq.enqueue_image("quick-start", img)
    time.sleep(3)
    print(outq.query("quick-start"))


if __name__ == "__main__":
    run()
