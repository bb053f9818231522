"""`pip install -e .` builds the two native libraries in-tree first (g++ and nvcc for sm_100a), the same
step `python __graft_entry__.py` performs (counterpart of the reference's build.sbt: one command, one build)."""
from setuptools import setup
from setuptools.command.build_py import build_py


class BuildNativeFirst(build_py):
    def run(self):
        import os
        import sys

        sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
        from akka_allreduce_b200 import build as native

        native.build_host()
        try:
            native.build_cuda()
        except RuntimeError as e:  # no nvcc: the CPU backend still works
            print("warning: CUDA library not built:", e)
        super().run()


setup(cmdclass={"build_py": BuildNativeFirst})
