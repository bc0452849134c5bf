"""Drop-in for the reference's ``utils`` module: put this repository root on
PYTHONPATH and ``from utils import *`` (ref:vae-hpo.py:16,
ref:example-subgroup.py:4) binds the B200 implementation.  The star-export
set equals the reference's (tests/golden/api.json: ``star_names``)."""
from multidisttorch_b200.utils import (  # noqa: F401
    dist, find_ifname, get_comm_size_and_rank, init_comm_size_and_rank, os, parse_slurm_nodelist, print0,
    psutil, re, setup_ddp, setup_ddp_groups, socket, torch,
)
