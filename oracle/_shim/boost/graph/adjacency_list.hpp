// TEST INFRASTRUCTURE ONLY -- a stand-in for the handful of boost::graph names that the reference's
// correct_links.cpp uses (correct_links.cpp:12-35,172-173,187-188,212-224,265,285-289), so that the UNMODIFIED
// reference source compiles into oracle/_ref/correct_links without boost (which is neither vendored nor installed).
// Nothing under salsa_b200/ includes this.  Semantics follow the documented behaviour of the boost calls:
//   adjacency_list<listS, vecS, directedS, ...> : vertices are 0..n-1, add_edge(u, v, w, g) grows the vertex set to
//       max(u, v) + 1, parallel edges are kept;
//   dijkstra_shortest_paths(g, s, distance_map(d).predecessor_map(p)) : every distance starts at
//       numeric_limits<Weight>::max(), every predecessor at the vertex itself, non-negative weights;
//   edge(u, v, g).second : is there an out-edge u -> v;
//   iterator_property_map<T*, IndexMap, T, T&> : p[index(v)].
#ifndef SALSA_ORACLE_BOOST_SHIM_ADJACENCY_LIST_HPP
#define SALSA_ORACLE_BOOST_SHIM_ADJACENCY_LIST_HPP
#include <climits>
#include <cstddef>
#include <limits>
#include <queue>
#include <sstream>
#include <string>
#include <utility>
#include <vector>

namespace boost {

struct listS {};
struct vecS {};
struct directedS {};
struct no_property {};
enum edge_weight_t { edge_weight };
enum vertex_name_t { vertex_name };
enum vertex_index_t { vertex_index };

template <class Tag, class T, class Base = no_property>
struct property {
    typedef T value_type;
    T value;
    property() : value() {}
    property(const T &v) : value(v) {}
};

template <class OutEdgeList, class VertexList, class Directed, class VertexProp, class EdgeProp>
class adjacency_list {
public:
    typedef std::size_t vertex_descriptor;
    typedef typename EdgeProp::value_type weight_type;
    struct out_edge {
        vertex_descriptor target;
        weight_type weight;
    };
    typedef std::pair<vertex_descriptor, std::size_t> edge_descriptor;      // (source, position in its out-edge list)
    std::vector<std::vector<out_edge> > out;
    std::size_t nedges;
    adjacency_list() : nedges(0) {}
};

template <class G>
struct graph_traits {
    typedef typename G::vertex_descriptor vertex_descriptor;
    typedef typename G::edge_descriptor edge_descriptor;
};

struct identity_index_map {
    std::size_t operator[](std::size_t v) const { return v; }
};
template <class G, class Tag>
struct property_map {
    typedef identity_index_map type;
};
template <class G>
inline identity_index_map get(vertex_index_t, const G &) { return identity_index_map(); }

template <class Iter, class IndexMap, class T, class Ref>
class iterator_property_map {
    Iter it;
    IndexMap idx;

public:
    iterator_property_map() : it(), idx() {}
    iterator_property_map(Iter i, IndexMap m) : it(i), idx(m) {}
    Ref operator[](std::size_t v) const { return it[idx[v]]; }
};

template <class O, class V, class D, class VP, class EP>
inline std::size_t num_vertices(const adjacency_list<O, V, D, VP, EP> &g) { return g.out.size(); }
template <class O, class V, class D, class VP, class EP>
inline std::size_t num_edges(const adjacency_list<O, V, D, VP, EP> &g) { return g.nedges; }

template <class O, class V, class D, class VP, class EP, class W>
inline std::pair<typename adjacency_list<O, V, D, VP, EP>::edge_descriptor, bool> add_edge(std::size_t u, std::size_t v, const W &w,
                                                                                            adjacency_list<O, V, D, VP, EP> &g)
{
    typedef adjacency_list<O, V, D, VP, EP> G;
    std::size_t need = (u > v ? u : v) + 1;
    if (g.out.size() < need) g.out.resize(need);
    typename G::out_edge e;
    e.target = v;
    e.weight = (typename G::weight_type)w;
    g.out[u].push_back(e);
    g.nedges++;
    return std::make_pair(typename G::edge_descriptor(u, g.out[u].size() - 1), true);
}

template <class O, class V, class D, class VP, class EP>
inline std::pair<typename adjacency_list<O, V, D, VP, EP>::edge_descriptor, bool> edge(std::size_t u, std::size_t v,
                                                                                        const adjacency_list<O, V, D, VP, EP> &g)
{
    typedef adjacency_list<O, V, D, VP, EP> G;
    if (u < g.out.size())
        for (std::size_t i = 0; i < g.out[u].size(); i++)
            if (g.out[u][i].target == v) return std::make_pair(typename G::edge_descriptor(u, i), true);
    return std::make_pair(typename G::edge_descriptor(0, 0), false);
}

// distance_map(d).predecessor_map(p)
template <class D, class P>
struct dijkstra_params {
    D dist;
    P pred;
};
template <class D>
struct dijkstra_params<D, void> {
    D dist;
    template <class P>
    dijkstra_params<D, P> predecessor_map(P p) const
    {
        dijkstra_params<D, P> r;
        r.dist = dist;
        r.pred = p;
        return r;
    }
};
template <class D>
inline dijkstra_params<D, void> distance_map(D d)
{
    dijkstra_params<D, void> r;
    r.dist = d;
    return r;
}

template <class G, class D, class P>
inline void dijkstra_shortest_paths(const G &g, std::size_t s, const dijkstra_params<D, P> &prm)
{
    typedef typename G::weight_type W;
    const std::size_t n = g.out.size();
    for (std::size_t v = 0; v < n; v++) {
        prm.dist[v] = std::numeric_limits<W>::max();
        prm.pred[v] = v;
    }
    if (s >= n) return;
    prm.dist[s] = W();
    typedef std::pair<W, std::size_t> item;
    std::priority_queue<item, std::vector<item>, std::greater<item> > pq;
    pq.push(item(W(), s));
    while (!pq.empty()) {
        item it = pq.top();
        pq.pop();
        const std::size_t u = it.second;
        if (it.first > prm.dist[u]) continue;
        for (std::size_t i = 0; i < g.out[u].size(); i++) {
            const std::size_t v = g.out[u][i].target;
            const W nd = prm.dist[u] + g.out[u][i].weight;
            if (nd < prm.dist[v]) {
                prm.dist[v] = nd;
                prm.pred[v] = u;
                pq.push(item(nd, v));
            }
        }
    }
}

}   // namespace boost
#endif
