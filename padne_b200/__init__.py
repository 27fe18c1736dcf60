"""padne_b200 -- B200-native implementation of padne's FEM hot path.

Scope: cotangent stiffness assembly, lumped/via/regulator coupling rows and
the sparse solve of ``padne/solver.py``, as hand-written sm_100a CUDA behind a C ABI
(``include/padne_b200.h``), with a Python host layer that mirrors the reference's
``solver.solve`` / ``assemble_system`` / ``solve_system`` seams.  There is no CPU fallback: every
compute entry point raises if the CUDA library is missing.
"""
__version__ = "0.1.0"
