"""CPU oracle for the GEMDAT hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import anything from this package, and there only
as the checker or as the timed CPU baseline.  ``gemdat_b200`` never imports it.

What is restated here (reference = GEMDAT v1.7.2 under /root/reference):

* ``oracle.states``   <- src/gemdat/transitions.py:736-809 (_calculate_atom_states)
* ``oracle.events``   <- src/gemdat/transitions.py:413-482 (_calculate_transition_events)
* ``oracle.fill``     <- src/gemdat/utils.py:59-125 (ffill / bfill / integer_remap)
* ``oracle.jumps``    <- src/gemdat/jumps.py:24-119, 185-221; transitions.py:812-834
* ``oracle.displ``    <- src/gemdat/trajectory.py:40-60,182-193,705-730; metrics.py:35-47,76-97,190-231
* ``oracle.msd``      <- src/gemdat/trajectory.py:896-933
* ``oracle.volume``   <- src/gemdat/volume.py:449-509
* ``oracle.radius``   <- src/gemdat/transitions.py:485-733 (SiteRadius)

Third-party arithmetic the reference reaches but does not vendor (so it is
restated from the published algorithms of the pinned versions, marked
[recalled] where no source was available in this container):

* ``oracle.mda``  MDAnalysis 2.6.1 (requirements_full.txt:57):
  lib.pkdtree.PeriodicKDTree, lib._augment, lib.distances.apply_PBC (ortho),
  lib.util.unique_rows.  The innermost ball query is the REAL scipy cKDTree.
* ``oracle.pmg``  pymatgen 2024.1.27 (requirements_full.txt:127):
  Lattice (matrix / parameters / metric tensor / get_all_distances) and
  core.trajectory.Trajectory.to_displacements / to_positions.

Pinning status (see DESIGN.md "Oracle"):
* pinned by the reference's own unit-test goldens (tests/trajectory_test.py,
  tests/metrics_test.py, tests/utils_test.py, tests/sites_test.py) and by
  fixtures produced by executing the reference's own modules in the build
  container with the third-party layer stubbed (tests/golden/make_golden.py);
* the real-data integration goldens need tests/data/short_simulation/vasprun.xml,
  a missing large blob: for _calculate_atom_states on real data: PARITY UNPINNED.
"""

NOSITE = -1
