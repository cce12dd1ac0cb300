"""Offline driver for maps with several object classes, with the reference's command line
(extract_map_objects_multi.py:35-98):
    python -m qsp_slam_b200.extract_map_objects_multi -y SETTINGS.yaml -m MAP_DIR [-n 128]
SETTINGS.yaml is the SLAM settings file (OpenCV FileStorage YAML: its first line ``%YAML:1.0`` is skipped, as the
reference does at :28-32); ``YoloClasses`` lists the class ids and ``DecoderPaths`` the DeepSDF experiment directory of
each.  Every object of MAP_DIR/MapObjects.txt is meshed by the extractor of ITS class id (second integer of its
first line) and written to MAP_DIR/objects/<id>.npy and <id>.ply."""
from __future__ import annotations

import argparse


def read_yaml_file(file_path):
    """extract_map_objects_multi.py:28-33: skip the ``%YAML:1.0`` line, parse the rest."""
    import yaml
    with open(file_path, "r") as f:
        f.readline()
        return yaml.load(f.read(), Loader=yaml.FullLoader)


def config_parser():
    parser = argparse.ArgumentParser()
    parser.add_argument('-y', '--yaml', type=str, required=True, help='path to yaml file')
    parser.add_argument('-m', '--map_dir', type=str, required=True, help='path to map directory')
    parser.add_argument('-n', '--voxels_dim', type=int, default=128, help='voxels resolution for running marching cube')
    return parser


def main(argv=None):
    from .deep_sdf.workspace import config_decoder
    from .reconstruct.optimizer import MeshExtractor
    from .map_io import extract_map_objects
    args = config_parser().parse_args(argv)
    yaml_data = read_yaml_file(args.yaml)
    yolo_classes, decoder_paths = yaml_data['YoloClasses'], yaml_data['DecoderPaths']
    code_len = 64
    extractors = {}
    for class_id, path in zip(yolo_classes, decoder_paths):
        extractors[int(class_id)] = MeshExtractor(config_decoder(path), code_len, args.voxels_dim)
    done = extract_map_objects(args.map_dir, extractors)
    print(f"extracted {len(done)} objects of {len(extractors)} classes into {args.map_dir}/objects")
    return done


if __name__ == "__main__":
    main()
