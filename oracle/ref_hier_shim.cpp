// TEST INFRASTRUCTURE ONLY — C entry points around the UNMODIFIED reference hierarchy file code
// (GH/hierarchy_loader.cpp, GH/hierarchy_writer.cpp, GH/traversal.cpp), compiled against oracle/eigen_stub (Eigen is used there
// only as POD containers). Built by oracle/Makefile into oracle/_ref/libref_hier.so; bound by oracle/refshim_hier.py.
#include "hierarchy_loader.h"
#include "hierarchy_writer.h"
#include "traversal.h"
#include <cstring>

namespace {
struct Loaded {
	std::vector<Eigen::Vector3f> pos; std::vector<SHs> shs; std::vector<float> alphas; std::vector<Eigen::Vector3f> scales;
	std::vector<Eigen::Vector4f> rot; std::vector<Node> nodes; std::vector<Box> boxes; std::vector<HierarchyNode> dnodes; int sh_degree = -1;
};
Loaded g;
}

extern "C" {
// static format: load into the shim's buffers, report sizes; ref_hier_fetch copies them out
int ref_hier_load(const char* fn, int* P, int* N) {
	try { g = Loaded(); HierarchyLoader::load(fn, g.pos, g.shs, g.alphas, g.scales, g.rot, g.nodes, g.boxes); }
	catch (const std::exception&) { return -1; }
	*P = (int)g.pos.size(); *N = (int)g.nodes.size(); return 0;
}
void ref_hier_fetch(float* pos, float* shs, float* alpha, float* scales, float* rot, int* nodes, float* boxes) {
	size_t P = g.pos.size(), N = g.nodes.size();
	memcpy(pos, g.pos.data(), P * 12); memcpy(shs, g.shs.data(), P * 192); memcpy(alpha, g.alphas.data(), P * 4);
	memcpy(scales, g.scales.data(), P * 12); memcpy(rot, g.rot.data(), P * 16);
	memcpy(nodes, g.nodes.data(), N * sizeof(Node)); memcpy(boxes, g.boxes.data(), N * sizeof(Box));
}
int ref_hier_write(const char* fn, int P, int N, float* pos, float* shs, float* alpha, float* scales, float* rot, int* nodes, float* boxes, int compressed) {
	try { HierarchyWriter w; w.write(fn, P, N, (Eigen::Vector3f*)pos, (SHs*)shs, alpha, (Eigen::Vector3f*)scales, (Eigen::Vector4f*)rot, (Node*)nodes, (Box*)boxes, compressed != 0); }
	catch (const std::exception&) { return -1; }
	return 0;
}
// dynamic format
int ref_dhier_load(const char* fn, int* P, int* sh_degree) {
	try { g = Loaded(); HierarchyLoader::loadDynamic(fn, g.pos, g.shs, g.alphas, g.scales, g.rot, g.dnodes, g.sh_degree); }
	catch (const std::exception&) { return -1; }
	*P = (int)g.pos.size(); *sh_degree = g.sh_degree; return 0;
}
// shs_floats = 3 * (deg + 1)^2: the reference reads the file's packed SH block into the front of its 48-float array
// (hierarchy_loader.cpp:160-175) and the torch binding reinterprets that front part (torch_interface.cpp:62-64)
void ref_dhier_fetch(float* pos, float* shs, int shs_floats, float* alpha, float* scales, float* rot, int* nodes) {
	size_t P = g.pos.size();
	memcpy(pos, g.pos.data(), P * 12); memcpy(shs, g.shs.data(), P * (size_t)shs_floats * 4); memcpy(alpha, g.alphas.data(), P * 4);
	memcpy(scales, g.scales.data(), P * 12); memcpy(rot, g.rot.data(), P * 16); memcpy(nodes, g.dnodes.data(), g.dnodes.size() * sizeof(HierarchyNode));
}
int ref_dhier_write(const char* fn, int P, int N, float* pos, float* shs, float* alpha, float* scales, float* rot, int* nodes, int sh_degree) {
	try { HierarchyWriter w; w.writeDynamic(fn, P, N, (Eigen::Vector3f*)pos, shs, alpha, (Eigen::Vector3f*)scales, (Eigen::Vector4f*)rot, (HierarchyNode*)nodes, sh_degree); }
	catch (const std::exception&) { return -1; }
	return 0;
}
// depth-target cut (GH/traversal.cpp:15-39)
long ref_expand_to_target(int* nodes, int target, int* out, long capacity) {
	std::vector<int> r = Traversal::expandToTarget((Node*)nodes, target);
	long n = (long)r.size();
	for (long i = 0; i < n && i < capacity; i++) out[i] = r[i];
	return n;
}
}
