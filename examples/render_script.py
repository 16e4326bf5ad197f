#!/usr/bin/env python
"""Headless counterpart of `pathtracer <scene.chai>` (reference src/main.cpp + sceneBuilder.h:313-329): load a scene script
(or a preset name: outside / sibenik / minecraft / 2mtris), restore the camera from a state file like save.txt, render
N samples and write a PFM.   usage: examples/render_script.py SCENE out.pfm [--state save.txt] [--spp 64] [--size 640x480]
[--assets DIR]: image files are looked up in DIR (default: the working directory, as the reference does) -- the MTL's map_Kd /
norm textures (scene.h:218-243), bluenoise.png and cave.hdr (pathtracer.h:81,127), decoded by cuda_pathtracer_b200/textures.py
with the reference loaders' conventions; a missing sky dome leaves the constant sky, a missing blue-noise tile the hash RNG."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from cuda_pathtracer_b200 import abi, api, scenes                                    # noqa: E402
from cuda_pathtracer_b200.scene_script import load_scene_script, read_state          # noqa: E402
from cuda_pathtracer_b200 import textures                                            # noqa: E402

PRESETS = {"outside": scenes.outside_scene, "sibenik": scenes.sibenik_scene, "minecraft": scenes.minecraft_scene, "2mtris": scenes.two_million_scene}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("scene"); ap.add_argument("out")
    ap.add_argument("--state", default="save.txt"); ap.add_argument("--spp", type=int, default=64); ap.add_argument("--size", default="640x480")
    ap.add_argument("--depth", type=int, default=32)                                # MAX_RAY_DEPTH, constants.h:12
    ap.add_argument("--assets", default="."); ap.add_argument("--sky", default="cave.hdr"); ap.add_argument("--blue-noise", default="bluenoise.png")
    a = ap.parse_args()
    w, h = (int(x) for x in a.size.split("x"))
    scene = PRESETS[a.scene]() if a.scene in PRESETS else load_scene_script(a.scene)   # getScene, sceneBuilder.h:313-329
    cam = read_state(a.state, w, h)                                                  # readState, stateLoader.h:54-82
    tex, nrm = textures.load_material_textures(scene, a.assets) if hasattr(scene, "textureName") else ({}, {})
    pt = api.Pathtracer(scene.arrays(), abi.default_config(w, h, max_ray_depth=a.depth, interactive_depth=scene.interactive_depth)).Init(textures=tex, normal_maps=nrm)
    sky, blue = os.path.join(a.assets, a.sky), os.path.join(a.assets, a.blue_noise)
    if os.path.exists(sky):
        pt.set_skydome(textures.load_texture_hdr(sky))                              # loadTextureHDR, pathtracer.h:81
    if os.path.exists(blue):
        pt.set_blue_noise(textures.load_texture_l(blue))                            # loadTextureL, pathtracer.h:127
    pt.Render(cam, shouldClear=True)
    for _ in range(a.spp):
        pt.Render(cam)
    pt.Finish()
    pt.write_pfm(a.out)
    print("wrote %s (%dx%d, %d spp)" % (a.out, w, h, a.spp))


if __name__ == "__main__":
    main()
