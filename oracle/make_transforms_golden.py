"""Test infrastructure: runs the UNMODIFIED reference parser (/root/reference/parser_instant_ngp.py) on the repo's real
robot capture (datasets/nerf-raw/wheel_real_robot: poses.txt from the forward kinematics + 12 frames) and commits its
output as fixtures for tests/test_poses_cpu.py:

    python oracle/make_transforms_golden.py          # needs /root/reference; writes tests/golden/wheel_transforms*.json

File paths are made relative and `sharpness` is dropped (it depends on the JPEG decoder build)."""
import json
import os
import subprocess
import sys
import tempfile

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def main():
    data = os.path.join(REF, "datasets", "nerf-raw", "wheel_real_robot")
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run([sys.executable, os.path.join(REF, "parser_instant_ngp.py"), "--poses", os.path.join(data, "poses.txt"),
                        "--image_folder", os.path.join(data, "images_robot"), "--recenter", "--output",
                        os.path.join(tmp, "transforms.json")], check=True, cwd=tmp, stdout=subprocess.DEVNULL)
        for suffix in ("", "_val", "_test", "_test_video"):
            d = json.load(open(os.path.join(tmp, "transforms%s.json" % suffix)))
            for fr in d["frames"]:
                fr.pop("sharpness", None)
                if "file_path" in fr:
                    fr["file_path"] = os.path.join("images_robot", os.path.basename(fr["file_path"]))
            with open(os.path.join(OUT, "wheel_transforms%s.json" % suffix), "w") as f:
                json.dump(d, f, indent=1)
            print("wrote wheel_transforms%s.json (%d frames)" % (suffix, len(d["frames"])))


if __name__ == "__main__":
    main()
