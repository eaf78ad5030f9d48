# coding=utf-8
"""JSON config -> constants.* (reference utils.py:9-32, same schema, no Taichi)."""
import json

import constants


def sys_init(config):
    constants.dim = config.get('dim')
    constants.auto_diff = 1 if config.get('auto_diff') else 0
    constants.use_explicit_method = config.get('use_explicit_method')
    constants.delta_time = config.get('delta_time')
    constants.g_dir = config.get('g_dir')
    constants.implicit_method = config.get('implicit_method')
    constants.preconditioned = config.get('preconditioned')
    constants.vec, constants.mat, constants.index = constants._make_types(constants.dim)
    # extension key (not in the reference's schema): which iterate order implicit_method 0 follows
    constants.jacobi_order = config.get('jacobi_order', 'simultaneous')


def read_config(file_name):
    try:
        with open(file_name, 'r') as f:
            return json.load(f)
    except Exception as e:  # same behaviour as the reference: print and exit(3)
        print(e)
        print('Parsing config file error')
        exit(3)
