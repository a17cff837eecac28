"""Packaging: same distribution name and console script as the reference (ref: setup.py:7-8,22-26).
The CUDA library is built in-tree by `python psana-ray_b200/build.py` (or __graft_entry__.build())."""
import setuptools

setuptools.setup(
    name="psana-ray",
    version="24.10.13+b200.1",
    description="psana-ray hot path (detector calibration + producer->consumer streaming) for NVIDIA B200",
    package_dir={"": "psana-ray_b200"},
    packages=["psana_ray"],
    entry_points={"console_scripts": ["psana-ray-producer=psana_ray.producer:main", "psana-ray-bridge=psana_ray.bridge:main"]},
    python_requires=">=3.9",
)
