"""Packaging stub (the reference ships one for its `trainer` package, /root/reference/setup.py:1-3).

The native libraries are built in-tree by `python -c "import __graft_entry__ as g; g.build()"`
(or `python -m distributed_tensorflow_b200.utils.build`) before packaging; they ride along as package data.
"""
from setuptools import find_packages, setup

setup(
    name="distributed_tensorflow_b200",
    version="1.0",
    description="B200-native parameter-server training framework with the interface of distributed_tensorflow",
    packages=find_packages(include=["trainer", "auto_stop_ps", "distributed_tensorflow_b200", "distributed_tensorflow_b200.*"]),
    package_data={"distributed_tensorflow_b200": ["_lib/*.so"]},
    python_requires=">=3.10",
)
