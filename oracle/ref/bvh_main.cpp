// Static-BVH harness: builds the reference's own BVH (enginee/src/ai/BVH.cpp, compiled unmodified) over a list of wall
// rectangles given in the order ai::init collects the statics (ai.cpp:28-36), and dumps the tree in pre-order, so that the
// product's host-side restatement (lc_static_bvh) can be pinned node for node.  TEST INFRASTRUCTURE ONLY.
//   stdin : int32 n_walls; float walls[n_walls][4] (origin x, z, width, height) in statics order
//   stdout: int32 n_nodes; per node (pre-order: node, left subtree, right subtree):
//           float aabb[4] (origin x, z, width, height; zeros for a leaf), int32 left, right (node indices; leaf: left =
//           position of the wall in the input list, right = -1)
#include "Pool.h"
#include "ai/BVH.h"
#include <cstdio>
#include <vector>

struct Node { float box[4]; int32_t left, right; };
static std::vector<Node> g_nodes;
static std::vector<Entity*> g_objs;

static int32_t walk(BVH* b) {
    const int32_t me = (int32_t)g_nodes.size();
    g_nodes.push_back(Node{{0, 0, 0, 0}, -1, -1});
    if (b->is_leaf()) {
        int32_t k = -1;
        for (size_t i = 0; i < g_objs.size(); ++i) if (g_objs[i] == b->o) k = (int32_t)i;
        g_nodes[me].left = k;
        return me;
    }
    g_nodes[me].box[0] = b->aabb._o.x; g_nodes[me].box[1] = b->aabb._o.y;
    g_nodes[me].box[2] = b->aabb._w; g_nodes[me].box[3] = b->aabb._h;
    const int32_t l = walk(b->left), r = walk(b->right);
    g_nodes[me].left = l; g_nodes[me].right = r;
    return me;
}

int main() {
    int32_t n = 0;
    if (fread(&n, 4, 1, stdin) != 1 || n < 1) return 2;
    std::vector<float> w(4 * (size_t)n);
    if (fread(w.data(), 4, w.size(), stdin) != w.size()) return 2;
    for (int32_t i = 0; i < n; ++i) {
        Entity& e = POOL.spawn_entity();
        uint16_t id = POOL.create<BoundVolume*>(new Rect(glm::vec2(w[4 * i], w[4 * i + 1]), w[4 * i + 2], w[4 * i + 3]));
        POOL.attach<BoundVolume*>(e, id);
        g_objs.push_back(&e);
    }
    BVH tree(g_objs);
    walk(&tree);
    int32_t m = (int32_t)g_nodes.size();
    fwrite(&m, 4, 1, stdout);
    fwrite(g_nodes.data(), sizeof(Node), g_nodes.size(), stdout);
    return 0;
}
