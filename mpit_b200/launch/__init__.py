"""Launchers (reference layer L6): ``run`` is the ``mpirun`` stand-in; claunch / glaunch / mlaunch /
plaunch / ptest / testreduceall mirror the reference scripts of the same names."""
