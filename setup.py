"""Packaging of mpit_b200 (the counterpart of the reference's rockspecs, /root/reference/mpit-openmpi-1.rockspec:15-24):
``python setup.py build_ext --inplace`` (or ``pip install --no-build-isolation -e .``) compiles the two native
libraries for sm_100a with ``tools/build.py`` and leaves them in ``mpit_b200/lib/``, where the package loads them from.
The libraries stay in the tree on purpose: nothing is JIT-compiled at run time and a GPU box receives them with the
sources."""
import os
import sys

from setuptools import Command, find_packages, setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))


def _build_native():
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import build as native_build          # tools/build.py
    native_build.build(verbose=True)


class BuildNative(Command):
    description = "compile libmpit_kernels.so (nvcc, sm_100a) and libmpit_host.so (g++) into mpit_b200/lib"
    user_options = [("inplace", "i", "accepted for compatibility: the libraries are always built in the tree")]
    boolean_options = ["inplace"]

    def initialize_options(self):
        self.inplace = True

    def finalize_options(self):
        pass

    def run(self):
        _build_native()


class BuildPyWithNative(build_py):
    def run(self):
        _build_native()
        super().run()


setup(
    name="mpit_b200",
    version="0.2.0",
    description="Asynchronous and elastic-averaging distributed SGD for 8xB200 NVSwitch boxes: message layer, parameter "
                "server, optimizers and tcgen05 kernels with the capabilities of fanshiqing/mpiT",
    packages=find_packages(include=["mpit_b200", "mpit_b200.*"]),
    package_data={"mpit_b200": ["lib/*.so"]},
    include_package_data=True,
    python_requires=">=3.10",
    install_requires=["torch", "numpy"],
    cmdclass={"build_ext": BuildNative, "build_native": BuildNative, "build_py": BuildPyWithNative},
    entry_points={"console_scripts": ["mpit-run=mpit_b200.launch.run:main"]},
)
