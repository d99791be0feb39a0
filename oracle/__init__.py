"""CPU oracle for the dipolar spin-ensemble time-evolution path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``quantumsimulations_b200/`` may import, call, link or execute anything in here.  Only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` use it, and only as the checker / timed CPU baseline.

What it restates (reference = /root/reference, files cited as DE = dipolar_ensemble_with_rare.py,
SW = sweep_sea_detuning.py):

* ``kron``      literal Kronecker-product build of H and the six observables       DE:37-52, DE:453-588
* ``dense``     dense/CSR H from a generic XYZ coupling table (independent of the product compiler)
* ``solvers``   X = exact eigendecomposition propagation; Q = SciPy zvode-Adams driven the way
                QuTiP 5's ``sesolve`` drives it (the reference's actual integrator)  DE:611-680
* ``metrics``   coarse-grain / slope / t-gated contrast                            SW:116-317
* ``c/``        plain-C matrix-free H-apply + embedded-RK port (the timed CPU baseline for large N)

Parity pin.  The reference ships no tests, fixtures or golden vectors, and its
arithmetic lives in QuTiP (version unpinned, README.md:71; dict-style options => QuTiP >= 5.0) which
is not installable here.  The pin used instead: ``ref_loader`` imports the reference's OWN
``dipolar_ensemble_with_rare.py`` from /root/reference with ``fake_qutip`` (a numpy/scipy stand-in
for the dozen QuTiP calls that file makes) so that the reference's Hamiltonian builder, initial
state and ``simulate_rare`` control flow execute unmodified; ``make_goldens.py`` stores what they
return under ``tests/golden/``.  The Hamiltonian/observable/initial-state half of the path is
therefore pinned by reference-generated fixtures; the integrator half (QuTiP -> zvode) is a
restatement against SciPy's own zvode and stays "parity unpinned" until a real QuTiP can be run.
"""
