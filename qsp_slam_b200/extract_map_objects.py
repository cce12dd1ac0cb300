"""Offline driver with the reference's command line (extract_map_objects.py:25-61):
    python -m qsp_slam_b200.extract_map_objects -c CONFIG.json -m MAP_DIR [-n 128]
reads MAP_DIR/MapObjects.txt and writes MAP_DIR/objects/<id>.npy and <id>.ply."""
from __future__ import annotations

import argparse


def config_parser():
    parser = argparse.ArgumentParser()
    parser.add_argument('-c', '--config', type=str, required=True, help='path to config file')
    parser.add_argument('-m', '--map_dir', type=str, required=True, help='path to map directory')
    parser.add_argument('-n', '--voxels_dim', type=int, default=128, help='voxels resolution for running marching cube')
    return parser


def main(argv=None):
    from .reconstruct.utils import get_configs
    from .reconstruct.utils import get_decoder
    from .reconstruct.optimizer import MeshExtractor
    from .map_io import extract_map_objects
    args = config_parser().parse_args(argv)
    configs = get_configs(args.config)
    decoder = get_decoder(configs)
    mesh_extractor = MeshExtractor(decoder, configs.optimizer.code_len, args.voxels_dim)
    done = extract_map_objects(args.map_dir, mesh_extractor)
    print(f"extracted {len(done)} objects into {args.map_dir}/objects")


if __name__ == "__main__":
    main()
