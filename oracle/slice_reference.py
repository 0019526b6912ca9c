"""TEST INFRASTRUCTURE — cuts the hot-path function bodies out of the reference's matcher / Frame / MapPoint sources.

The reference's ORBmatcher.cc, Frame.cc and MapPoint.cc cannot be compiled whole (they pull the entire SLAM graph, DBoW2
and g2o), but the functions SURVEY.md section 8 names are self-contained once the handful of members they touch exists.
This script copies those functions, BY PATH AND LINE RANGE, from the reference tree into git-ignored
oracle/_ref/gen_*.inc files at build time; oracle/refmatch/refmatch.cpp then compiles them, unmodified, between stub
class declarations (oracle/refmatch/stubs.hpp).  Nothing of the reference is stored in this repository.

Every range is checked: its first line must start with the expected signature and its last line must be the closing
brace, so a reference tree with different line numbers fails the build instead of silently compiling something else.

  python oracle/slice_reference.py /root/reference oracle/_ref
"""
import os
import sys

SLICES = {
    "gen_frame_defs.inc": ("include/Frame.h", [
        (36, 37, "#define FRAME_GRID_ROWS", None),
    ]),
    "gen_orbmatcher.inc": ("src/ORBmatcher.cc", [
        (37, 43, "const int ORBmatcher::TH_HIGH", "}"),
        (45, 124, "int ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints", "}"),
        (126, 132, "float ORBmatcher::RadiusByViewingCos", "}"),
        (400, 515, "int ORBmatcher::SearchForInitialization(", "}"),
        (1292, 1423, "int ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame", "}"),
        (1554, 1595, "void ORBmatcher::ComputeThreeMaxima(", "}"),
        (1605, 1621, "int ORBmatcher::DescriptorDistance(", "}"),
    ]),
    "gen_frame.inc": ("src/Frame.cc", [
        (114, 129, "void Frame::AssignFeaturesToGrid()", "}"),
        (209, 262, "vector<size_t> Frame::GetFeaturesInArea(", "}"),
        (264, 274, "bool Frame::PosInGrid(", "}"),
        (286, 320, "void Frame::UndistortKeyPoints()", "}"),
        (322, 353, "void Frame::ComputeImageBounds(", "}"),
        (355, 384, "void Frame::antidistorsionarProyeccionEquidistante(", "}"),
    ]),
    "gen_mappoint.inc": ("src/MapPoint.cc", [
        (227, 292, "void MapPoint::ComputeDistinctiveDescriptors()", "}"),
    ]),
    "gen_osmap.inc": ("src/Osmap.cpp", [
        (775, 779, "void Osmap::serialize(const Mat &m, SerializedDescriptor", "}"),
        (781, 786, "void Osmap::deserialize(const SerializedDescriptor", "}"),
        (819, 824, "void Osmap::serialize(const KeyPoint &kp, SerializedKeypoint", "}"),
        (826, 831, "void Osmap::deserialize(const SerializedKeypoint", "}"),
    ]),
}


def main(ref, out):
    os.makedirs(out, exist_ok=True)
    for name, (rel, ranges) in SLICES.items():
        path = os.path.join(ref, rel)
        lines = open(path, encoding="utf-8", errors="replace").read().split("\n")
        chunks = []
        for a, b, first, last in ranges:
            body = lines[a - 1:b]
            if not body[0].startswith(first) or (last is not None and body[-1].rstrip() != last):
                sys.exit("slice_reference: %s:%d-%d does not hold the expected function (%r ... %r)" % (path, a, b, body[0][:60], body[-1][:20]))
            chunks.append('#line %d "%s"\n' % (a, path) + "\n".join(body) + "\n")
        with open(os.path.join(out, name), "w", encoding="utf-8") as f:
            f.write("// GENERATED at build time from %s (line ranges %s) — git-ignored, never committed.\n" % (
                path, ", ".join("%d-%d" % (a, b) for a, b, _, _ in ranges)))
            f.write("\n".join(chunks))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
