"""Generates tests/golden/fk_tree.json by PARSING the reference's MuJoCo model file
/root/reference/diffusion/assets/dp_env_v2.xml (the model mocap_player.py:12,67-73 loads): for every body its parent,
its offset in the parent frame, its joints in document order (qpos index, type, axis, anchor) and its geoms (mass,
centre in the body frame; a capsule's centre is the midpoint of its `fromto`).  Build container only:
    python tests/golden/make_fk_tree.py
tests/test_fk_tree.py checks the hand-written tables of oracle/scoring_oracle.py and the arithmetic of
csrc/fk_math.cuh against this file, so a transcription error in either shows up as a test failure."""
import json
import os
import xml.etree.ElementTree as ET

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/diffusion/assets/dp_env_v2.xml"


def vec(s):
    return [float(v) for v in s.split()]


def main():
    root = ET.parse(SRC).getroot()
    compiler = root.find("compiler").attrib
    bodies, qpos_next = [], [0]

    def walk(elem, parent):
        name = elem.attrib["name"]
        joints = []
        for j in elem.findall("joint"):
            jt = j.attrib.get("type", "hinge")
            n = 7 if jt == "free" else 1
            joints.append({"name": j.attrib["name"], "type": jt, "qpos": qpos_next[0],
                           "axis": vec(j.attrib["axis"]) if "axis" in j.attrib else None,
                           "pos": vec(j.attrib.get("pos", "0 0 0")),
                           "range": vec(j.attrib["range"]) if "range" in j.attrib else None})
            qpos_next[0] += n
        geoms = []
        for g in elem.findall("geom"):
            if "fromto" in g.attrib:
                ft = vec(g.attrib["fromto"])
                centre = [(ft[i] + ft[i + 3]) / 2 for i in range(3)]
            else:
                centre = vec(g.attrib.get("pos", "0 0 0"))
            geoms.append({"name": g.attrib["name"], "mass": float(g.attrib["mass"]), "centre": centre,
                          "type": g.attrib.get("type", "sphere")})
        bodies.append({"name": name, "parent": parent, "pos": vec(elem.attrib.get("pos", "0 0 0")), "joints": joints,
                       "geoms": geoms})
        for child in elem.findall("body"):
            walk(child, name)

    for b in root.find("worldbody").findall("body"):
        walk(b, None)
    out = {"source": "diffusion/assets/dp_env_v2.xml", "compiler": compiler, "nq": qpos_next[0], "bodies": bodies}
    json.dump(out, open(os.path.join(HERE, "fk_tree.json"), "w"), indent=1)
    print("bodies", len(bodies), "nq", qpos_next[0], "total mass", sum(g["mass"] for b in bodies for g in b["geoms"]))


if __name__ == "__main__":
    main()
