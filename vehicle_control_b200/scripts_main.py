"""Headless equivalent of vehicle_control/scripts/main.py:1-14: load the three YAMLs, build the simulator, run, summarize.

    python -m vehicle_control_b200.scripts_main [simconfig.yaml]
"""
import os
import sys

from .simulation.racing import RacingSimulator
from .utils.common_utils import config_dir, load_config


def main(simconfig_path=None):
    cd = config_dir()
    simconfig = load_config(simconfig_path or os.path.join(cd, "simconfig.yaml"))
    trackconfig = load_config(os.path.join(cd, "environment", f"{simconfig.track_name}.yaml"))
    carconfig = load_config(os.path.join(cd, "models", "dynamic_car.yaml"))
    sim = RacingSimulator(simconfig, carconfig, trackconfig)
    sim.run()
    sim.summarize()
    return sim


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else None)
