"""Copies the reference's own problem data file (alpaqa/test/outer/rosenbrock_functions_test.csv, loaded by
alpaqa/test/outer/test-casadi.cpp:31-43) into tests/golden/ as the known-answer input of the data-file reader.
Run in the build container (needs /root/reference)."""
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
shutil.copyfile("/root/reference/alpaqa/test/outer/rosenbrock_functions_test.csv",
                os.path.join(HERE, "rosenbrock_functions_test.csv"))
