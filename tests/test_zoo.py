"""zoo.py generates the same graphs as the reference YAML files."""
import pytest
from conftest import HAVE_REF

from yolo_series_b200 import zoo


def test_zoo_shapes():
    n = {k: len(zoo.get_cfg(k)['backbone']) + len(zoo.get_cfg(k)['head']) for k in zoo.ZOO}
    assert n == {'yolov7': 106, 'yolov7-tiny': 78, 'yolov7-w6': 123, 'yolov7-w6-custom': 123, 'yolov7-e6e': 266}
    assert zoo.get_cfg('yolov7-w6-custom')['nc'] == 83


@pytest.mark.skipif(not HAVE_REF, reason='/root/reference not mounted')
@pytest.mark.parametrize('name', sorted(zoo.ZOO))
def test_zoo_equals_reference_yaml(name):
    import yaml
    with open(f'/root/reference/cfg/training/{name}.yaml') as f:
        ref = yaml.safe_load(f)
    mine = zoo.get_cfg(name)
    for k in ('nc', 'depth_multiple', 'width_multiple', 'anchors', 'backbone', 'head'):
        assert ref[k] == mine[k], k
