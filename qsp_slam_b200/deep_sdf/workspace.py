"""config_decoder (deep_sdf/workspace.py:202-224): reads ``specs.json`` + ``ModelParameters/latest.pth``
and returns a device-resident decoder.  The checkpoint format is unchanged: ``model_state_dict`` with
DataParallel's ``module.`` key prefix."""
from __future__ import annotations

import json
import os

from ..decoder import DeviceDecoder

model_params_subdir = "ModelParameters"


def config_decoder(experiment_directory, checkpoint="latest", device: int = 0):
    specs_filename = os.path.join(experiment_directory, "specs.json")
    if not os.path.isfile(specs_filename):
        raise Exception('The experiment directory does not include specifications file "specs.json"')
    specs = json.load(open(specs_filename))
    net = specs.get("NetworkSpecs", {})
    if specs.get("CodeLength", 64) != 64 or list(net.get("dims", [512] * 8)) != [512] * 8 or \
            list(net.get("latent_in", [4])) != [4] or not net.get("weight_norm", True) or net.get("xyz_in_all", False):
        raise ValueError("only the standard DeepSDF layout is supported: code 64, 8x512, latent_in=[4], weight norm")
    import torch
    saved = torch.load(os.path.join(experiment_directory, model_params_subdir, checkpoint + ".pth"),
                       map_location="cpu")
    return DeviceDecoder(saved["model_state_dict"], device)
