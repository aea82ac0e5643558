"""Minimal ``mpi4py`` stand-in used ONLY to run the unmodified reference (harness, not product).

The reference imports ``from mpi4py import MPI`` in a class body
(``mpitree/tree/decision_tree.py:313``) and uses ``MPI.COMM_WORLD``, ``Get_rank``,
``Get_size``, ``Split(color, key)``, the pickled lower-case ``allgather`` and
``Free`` — nothing else.  Real mpi4py / an MPI runtime are not installable in
this image (no network, no wheel), so this shim provides those six calls over
``torch.distributed`` (gloo) when launched by torchrun, and a serial world
otherwise.  Numbers produced through it are labelled "reference code over shim
transport".
"""
