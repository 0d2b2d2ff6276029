/*
 * host_demo.cpp — TEST PROGRAM for include/rt3_host.hpp: a C++ host that links librt3.so directly (no Python, no torch) and drives
 * the path through the reference-shaped builder API.  tests/test_cpp_host.py writes a scene + a list of GConst blocks into a blob,
 * runs this program and compares what it writes with the CPU oracle, bit for bit.
 *
 *   host_demo <scene blob> <output file>
 *
 * blob:   "RT3S", u32 W, H, n_vertices, n_indices, n_geometries, n_frames, env_w, env_h;
 *         rt3_vertex[n_vertices], u32[n_indices], rt3_geometry_info[n_geometries], float[env_w * env_h * 3], rt3_params[n_frames]
 * output: Light (RGBA32F) of every frame, then gbuffer (RGBA32UI), depth (R32F) and the post-processed image (RGBA32F) of the last
 *         frame, then rays_primary, rays_secondary, paths_started (u64 each)
 * exit:   0 ok, 3 no CUDA device (rt3_create refused: there is no CPU fallback), 1 anything else
 */
#include <cstdio>
#include <cstdlib>
#include <fstream>

#include "rt3_host.hpp"

template <class T>
static std::vector<T> read_array(std::ifstream& f, size_t n) {
    std::vector<T> v(n);
    f.read((char*)v.data(), (std::streamsize)(n * sizeof(T)));
    if (!f) throw std::runtime_error("scene blob is truncated");
    return v;
}

int main(int argc, char** argv) {
    if (argc != 3) { fprintf(stderr, "usage: host_demo <scene blob> <output file>\n"); return 1; }
    try {
        std::ifstream in(argv[1], std::ios::binary);
        char magic[4];
        in.read(magic, 4);
        if (!in || memcmp(magic, "RT3S", 4) != 0) throw std::runtime_error("not a scene blob");
        const std::vector<uint32_t> h = read_array<uint32_t>(in, 8);
        const uint32_t W = h[0], H = h[1], nv = h[2], ni = h[3], ng = h[4], nf = h[5], ew = h[6], eh = h[7];
        const auto vertices = read_array<rt3_vertex>(in, nv);
        const auto indices = read_array<uint32_t>(in, ni);
        const auto geometries = read_array<rt3_geometry_info>(in, ng);
        const auto env = read_array<float>(in, (size_t)ew * eh * 3);
        const auto frames = read_array<rt3_params>(in, nf);

        rt3h::Context ctx(0);
        rt3h::RenderWorld world(ctx);
        world.push_vertices(vertices.data(), nv);                  /* loaded_assets, world/mod.rs:83-101 */
        world.push_indices(indices.data(), ni);
        for (const rt3_geometry_info& g : geometries) world.push_geometry(g);
        if (ew) world.set_environment(ew, eh, env.data());
        const rt3_build_stats bs = world.create_acceleration_structure();
        fprintf(stderr, "host_demo: %u triangles, %u wide nodes, build %.3f ms\n", bs.triangles, bs.wide_nodes, bs.ms_total);

        /* error behaviour of the builder: an unbound slot and an unknown shader are reported, not ignored */
        rt3h::Renderer r(world, W, H);
        int caught = 0;
        try { rt3h::RayTracingPass::create(r.rg, "bad").shader("gbuffer").constants(frames[0]).write("gbuffer", "gbuffer").launch(); } catch (const rt3h::Error&) { caught++; }
        try { rt3h::RayTracingPass::create(r.rg, "bad").shader("no_such_shader").constants(frames[0]).launch(); } catch (const rt3h::Error&) { caught++; }
        try { rt3h::ComputePass::create(r.rg, "bad").shader("postprocess").read("Depth", "gbuffer_depth").dispatch(); } catch (const rt3h::Error&) { caught++; }
        if (caught != 3) throw std::runtime_error("builder errors were not reported");

        std::ofstream out(argv[2], std::ios::binary);
        const size_t px = (size_t)W * H;
        std::vector<float> light(px * 4), depth(px), post(px * 4);
        std::vector<uint32_t> gbuffer(px * 4);
        ctx.counters(true);
        for (uint32_t i = 0; i < nf; ++i) {
            const std::string name = r.frame(frames[i], i + 1 == nf);
            r.rg.download(name, light.data());
            out.write((const char*)light.data(), (std::streamsize)(light.size() * 4));
        }
        r.rg.download("gbuffer", gbuffer.data());
        r.rg.download("gbuffer_depth", depth.data());
        r.rg.download("out", post.data());
        out.write((const char*)gbuffer.data(), (std::streamsize)(gbuffer.size() * 4));
        out.write((const char*)depth.data(), (std::streamsize)(depth.size() * 4));
        out.write((const char*)post.data(), (std::streamsize)(post.size() * 4));
        const rt3_counters c = ctx.counters();
        const uint64_t tail[3] = {c.rays_primary, c.rays_secondary, c.paths_started};
        out.write((const char*)tail, sizeof(tail));
        if (!out) throw std::runtime_error("cannot write the output file");
        return 0;
    } catch (const rt3h::Error& e) {
        fprintf(stderr, "host_demo: rt3 error %d: %s\n", e.status, e.what());
        return e.status == RT3_ERR_NO_DEVICE ? 3 : 1;
    } catch (const std::exception& e) {
        fprintf(stderr, "host_demo: %s\n", e.what());
        return 1;
    }
}
