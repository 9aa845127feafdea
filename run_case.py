#!/usr/bin/env python
"""Driver with the reference's command line (src/run_case.py:14-26,152-185):

    python run_case.py -case cavity -ksp_type cg -pc_type jacobi -ksp_rtol 1e-8 -ts_adapt_type none -ts_dt 0.01 \
                       [-log INFO] [-nelem 64] [-ngl 5] [-ts_max_steps 10]

`-case` picks the class and `pynama_b200/cases/<case>.yaml`; every `-ksp_* / -pc_* / -ts_*` option is read by the
solvers' setFromOptions.  The chart / ParaView studies of the reference's `-test` switch are plotting code and are not
part of the hot path."""
import logging
import os
import sys

import yaml

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from pynama_b200.common.options import Options  # noqa: E402

Options.init(sys.argv)
OptDB = Options()
case = OptDB.getString('case', False)
customFunctions = ['taylor-green']

if case in customFunctions:
    from pynama_b200.cases.custom_func import CustomFuncCase as FemProblem
elif case == 'uniform':
    from pynama_b200.cases.uniform import UniformFlow as FemProblem
elif case == 'cavity':
    from pynama_b200.cases.cavity import Cavity as FemProblem
elif case == 'ibm-static':
    from pynama_b200.cases.immersed_boundary import ImmersedBoundaryStatic as FemProblem
elif case == 'ibm-dynamic':
    from pynama_b200.cases.immersed_boundary import ImmersedBoundaryDynamic as FemProblem
else:
    print("Case not defined unabled to import")
    sys.exit()


def timeSolving(config):
    fem = FemProblem(config)
    fem.setUp()
    fem.setUpSolver()
    if not fem.comm.rank:
        fem.logger.info("Solving problem...")
    fem.timer.tic()
    fem.startSolver()
    if not fem.comm.rank:
        fem.logger.info(f"Solver Finished in {fem.timer.toc()} seconds")
        fem.logger.info(f"Total time: {fem.timerTotal.toc()} seconds")
    return fem


def main():
    log = OptDB.getString('log', 'INFO')
    logging.basicConfig(level=log.upper())
    logger = logging.getLogger("Init")
    path = os.path.join(ROOT, "pynama_b200", "cases", f"{case}.yaml")
    try:
        with open(path) as f:
            yamlData = yaml.load(f, Loader=yaml.Loader)
    except OSError:
        logger.info(f"Case '{case}' Not Found")
        return
    # convenience overrides of the YAML (the reference edits the file instead)
    if OptDB.hasName('nelem'):
        n = OptDB.getInt('nelem')
        yamlData['domain']['box-mesh']['nelem'] = [n, n]
    if OptDB.hasName('ngl'):
        yamlData['domain']['ngl'] = OptDB.getInt('ngl')
    timeSolving(yamlData)


if __name__ == "__main__":
    main()
