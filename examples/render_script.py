#!/usr/bin/env python
"""Headless counterpart of `pathtracer <scene.chai>` (reference src/main.cpp + sceneBuilder.h:313-329): load a scene script
(or a preset name: outside / sibenik / minecraft / 2mtris), restore the camera from a state file like save.txt, render
N samples and write a PFM.   usage: examples/render_script.py SCENE out.pfm [--state save.txt] [--spp 64] [--size 640x480]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from cuda_pathtracer_b200 import abi, api, scenes                                    # noqa: E402
from cuda_pathtracer_b200.scene_script import load_scene_script, read_state          # noqa: E402

PRESETS = {"outside": scenes.outside_scene, "sibenik": scenes.sibenik_scene, "minecraft": scenes.minecraft_scene, "2mtris": scenes.two_million_scene}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("scene"); ap.add_argument("out")
    ap.add_argument("--state", default="save.txt"); ap.add_argument("--spp", type=int, default=64); ap.add_argument("--size", default="640x480")
    ap.add_argument("--depth", type=int, default=32)                                # MAX_RAY_DEPTH, constants.h:12
    a = ap.parse_args()
    w, h = (int(x) for x in a.size.split("x"))
    scene = PRESETS[a.scene]() if a.scene in PRESETS else load_scene_script(a.scene)   # getScene, sceneBuilder.h:313-329
    cam = read_state(a.state, w, h)                                                  # readState, stateLoader.h:54-82
    pt = api.Pathtracer(scene.arrays(), abi.default_config(w, h, max_ray_depth=a.depth, interactive_depth=scene.interactive_depth)).Init()
    pt.Render(cam, shouldClear=True)
    for _ in range(a.spp):
        pt.Render(cam)
    pt.Finish()
    pt.write_pfm(a.out)
    print("wrote %s (%dx%d, %d spp)" % (a.out, w, h, a.spp))


if __name__ == "__main__":
    main()
