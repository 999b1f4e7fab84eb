"""Micro-benchmarks shipped with the package (role of the reference's ``benchmarks/`` tree used by ``bin/ds_bench``)."""
