// ORACLE — TEST INFRASTRUCTURE ONLY.
// Fixture generator: parses an OBJ file with the oracle's restatement of the reference's
// record grammar (src/obj.rs:53-184) and writes the binary mesh fixture that scenes.hpp's
// load_mesh_fixture() reads (the GPU box has no /root/reference, so the mesh travels as data/).
//   usage: obj2mesh <in.obj> <out.mesh>
#include "rt_oracle.hpp"

int main(int argc, char** argv) {
    if (argc != 3) { std::fprintf(stderr, "usage: %s in.obj out.mesh\n", argv[0]); return 2; }
    std::ifstream in(argv[1]);
    if (!in) { std::fprintf(stderr, "cannot open %s\n", argv[1]); return 1; }
    orc::ObjData d = orc::ObjParser::parse_text(in);
    std::ofstream out(argv[2], std::ios::binary);
    out.write("RTCMESH1", 8);
    uint32_t hdr[4] = {(uint32_t)d.vertices.size(), (uint32_t)d.normals.size(), (uint32_t)d.triangles.size(), (uint32_t)d.groups.size()};
    out.write((const char*)hdr, sizeof hdr);
    for (auto& v : d.vertices) { double p[3] = {v.x, v.y, v.z}; out.write((const char*)p, 24); }
    for (auto& v : d.normals) { double p[3] = {v.x, v.y, v.z}; out.write((const char*)p, 24); }
    for (auto& t : d.triangles) {
        uint32_t r[7] = {(uint32_t)t.p[0], (uint32_t)t.p[1], (uint32_t)t.p[2], t.has_n ? (uint32_t)t.n[0] : 0u,
                         t.has_n ? (uint32_t)t.n[1] : 0u, t.has_n ? (uint32_t)t.n[2] : 0u, (uint32_t)t.group};
        out.write((const char*)r, sizeof r);
    }
    for (auto& g : d.groups) { uint32_t len = (uint32_t)g.size(); out.write((const char*)&len, 4); out.write(g.data(), len); }
    std::printf("%s: %zu v, %zu vn, %zu triangles, %zu groups, %zu ignored lines\n", argv[2], d.vertices.size(), d.normals.size(),
                d.triangles.size(), d.groups.size(), d.ignored.size());
    return 0;
}
