"""Packaging (role of the reference's ``setup.py``: install the package, pre-build the native ops, install the CLI
entry points).

    pip install --no-build-isolation .          # builds ops/lib/libhds_b200.so (nvcc, sm_100a) + libhds_host.so (g++)
    HDS_BUILD_OPS=0 pip install .               # python only; ops are JIT-built on first use (ops/op_builder)
    python setup.py build_ops                   # just (re)build the in-tree libraries

The native libraries are plain C-ABI shared objects loaded with ctypes, so there are no torch / pybind extension
modules to declare: ``build_py`` calls the in-tree build (``hcache_deepspeed_b200/ops/build.py``) and ships the result as
package data together with the CUDA / C++ sources (needed for JIT builds of single ops).
"""
import importlib.util
import os
import sys

from setuptools import Command, find_packages, setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = "hcache_deepspeed_b200"


def _load(relpath, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, relpath))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def build_native(force=False):
    """Compile both in-tree libraries; returns False (with the reason printed) when the tool chain is missing."""
    if os.environ.get("HDS_BUILD_OPS", "1") == "0":
        print("[setup] HDS_BUILD_OPS=0: skipping the native build (ops will be JIT-built on first use)")
        return False
    sys.path.insert(0, ROOT)
    try:
        from hcache_deepspeed_b200.ops import build
        build.build_host(force=force, verbose=True)
        try:
            build.build_cuda(force=force, verbose=True)
        except RuntimeError as e:       # no nvcc: a CPU-only install still works (torch back end + host library)
            print(f"[setup] CUDA library not built: {e}")
            return False
        return True
    finally:
        sys.path.pop(0)


class BuildOps(Command):
    description = "build the in-tree native libraries (sm_100a kernels + host runtime)"
    user_options = [("force", "f", "rebuild even if up to date")]

    def initialize_options(self):
        self.force = False

    def finalize_options(self):
        self.force = bool(self.force)

    def run(self):
        build_native(force=self.force)


class BuildPyWithOps(build_py):

    def run(self):
        build_native()
        super().run()
        # the public headers travel inside the package too: JIT builds of single ops in an installed tree need them
        import shutil
        dst = os.path.join(self.build_lib, PKG, "include")
        os.makedirs(dst, exist_ok=True)
        for h in ("hds_b200.h", "hds_host.h"):
            shutil.copy2(os.path.join(ROOT, "include", h), os.path.join(dst, h))


version = _load(os.path.join(PKG, "version.py"), "_hds_version").__version__

setup(
    name="hcache-deepspeed-b200",
    version=version,
    description="B200-native (sm_100a) LLM serving and training framework with hidden-state cache (HCache) restore",
    long_description=open(os.path.join(ROOT, "README.md")).read(),
    long_description_content_type="text/markdown",
    packages=find_packages(include=[PKG, PKG + ".*"]),
    package_data={PKG: ["ops/csrc/*.cu", "ops/csrc/*.cuh", "ops/csrc_host/*.cpp", "ops/csrc_host/*.h", "ops/lib/*.so",
                        ]},
    data_files=[("include", ["include/hds_b200.h", "include/hds_host.h"])],
    scripts=["bin/hds", "bin/hds_report", "bin/hds_bench", "bin/hds_elastic", "bin/hds_ssh", "bin/hds_io", "bin/hds_nvme_tune"],
    python_requires=">=3.10",
    install_requires=["torch>=2.6", "numpy", "pydantic>=2", "packaging", "psutil", "tqdm"],
    extras_require={"hf": ["transformers", "safetensors"], "dev": ["pytest", "pytest-timeout"]},
    cmdclass={"build_py": BuildPyWithOps, "build_ops": BuildOps},
    zip_safe=False,
)
