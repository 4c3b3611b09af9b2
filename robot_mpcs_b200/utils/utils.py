"""Host helpers mirroring robotmpcs/utils/utils.py (only the parts on the MPC-solve path)."""
import yaml


def parse_setup(setup_file: str) -> dict:
    """YAML config -> dict {mpc, robot, example}  (reference: robotmpcs/utils/utils.py:5-8)."""
    with open(setup_file, "r") as stream:
        return yaml.safe_load(stream)
