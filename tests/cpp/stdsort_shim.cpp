// The REAL libstdc++ std::sort with the reference's comparator (src/ORBextractor.cc:209-224,
// called at :373), exposed to tests/cv2_compose.py through ctypes.  The comparator is not a total
// order on the nodes (ties on (size, UL.x) with different UL.y), so the order of tied elements is
// whatever introsort leaves -- the Python re-composition of the extractor must not guess it.
// Test infrastructure only.
#include <algorithm>
#include <utility>
#include <vector>

namespace {
struct Node { int ulx; int id; };
bool compareNodes(std::pair<int, Node*>& e1, std::pair<int, Node*>& e2) {
    if (e1.first < e2.first) return true;
    else if (e1.first > e2.first) return false;
    else return e1.second->ulx < e2.second->ulx;
}
}  // namespace

extern "C" void sort_nodes(int n, const int* size, const int* ulx, int* order_out) {
    std::vector<Node> nodes(n);
    std::vector<std::pair<int, Node*>> v;
    v.reserve(n);
    for (int i = 0; i < n; ++i) { nodes[i].ulx = ulx[i]; nodes[i].id = i; v.push_back(std::make_pair(size[i], &nodes[i])); }
    std::sort(v.begin(), v.end(), compareNodes);
    for (int i = 0; i < n; ++i) order_out[i] = v[i].second->id;
}
