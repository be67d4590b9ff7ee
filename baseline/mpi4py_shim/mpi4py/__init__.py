"""Stand-in for the ``mpi4py`` package, just large enough to run the reference script unmodified
on a box with no MPI runtime.  Transport: ``torch.distributed`` gloo.  This is a
measurement harness for the baseline, not part of the b200dp product."""
__version__ = "0.0-gloo-shim"
