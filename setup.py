"""Packaging (reference: pyzoo/setup.py packs the jar + scripts; here the native libraries are built in-tree first)."""
import os
import subprocess
import sys

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py

HERE = os.path.dirname(os.path.abspath(__file__))


class BuildWithNative(build_py):
    def run(self):
        subprocess.check_call([sys.executable, "-m", "zoo.native.build"], cwd=HERE)
        super().run()


setup(name="zoo-b200", version="0.8.0.dev0", description="Blackwell-native framework with the capabilities of Analytics Zoo",
      packages=find_packages(include=["zoo", "zoo.*"]), package_data={"zoo.native": ["*.so"]}, python_requires=">=3.10",
      install_requires=["torch>=2.6", "numpy", "pandas", "pyyaml", "cloudpickle", "scipy"],
      extras_require={"image": ["opencv-python", "torchvision"], "automl": ["scikit-learn"]},
      scripts=[os.path.join("scripts", s) for s in ("zoo-submit", "jupyter-with-zoo.sh")] +
              [os.path.join("scripts", "cluster-serving", s) for s in os.listdir(os.path.join(HERE, "scripts", "cluster-serving"))
               if s.startswith("cluster-serving-")],
      cmdclass={"build_py": BuildWithNative})
