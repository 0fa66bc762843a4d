"""CPU oracle for the catenary / RK4 / sampling-MPC hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import this
directory; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs do, and only as the checker or the
timed CPU arm.  See ``oracle/catenary_oracle.py`` for the parity status of each
function.
"""
