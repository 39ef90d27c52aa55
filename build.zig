// build.zig - the reference drives its build from zig by shelling out to the compiler (reference build.zig:48-93: one
// g++ command per binary).  This file keeps that style: `zig build` runs CMake + Ninja on CMakeLists.txt, which compiles the
// CUDA library for sm_100a (nvcc) and the host binaries (g++), and installs nothing - the artefacts land where `make` puts
// them (deepreadmapper_b200/lib/libdrm_cuda.so, build/bin/{pipeline,hnswpq_index,test_sw_score,test_search}).
//
// UNVERIFIED: zig is not installed in the build image (no network), so this file has never been run; CMakeLists.txt is the
// build that __graft_entry__.build() exercises.  Steps: `zig build` (configure + build), `zig build pipeline`,
// `zig build hnswpq_index`.
const std = @import("std");

pub fn build(b: *std.Build) void {
    const build_dir = "build/cmake";

    const configure = b.addSystemCommand(&[_][]const u8{
        "cmake", "-S", ".", "-B", build_dir, "-G", "Ninja", "-DCMAKE_BUILD_TYPE=Release", "-DCMAKE_CUDA_ARCHITECTURES=100a",
    });

    const all = b.addSystemCommand(&[_][]const u8{ "ninja", "-C", build_dir });
    all.step.dependOn(&configure.step);
    b.getInstallStep().dependOn(&all.step);

    const targets = [_][]const u8{ "drm_cuda", "pipeline", "hnswpq_index", "test_sw_score", "test_search" };
    for (targets) |t| {
        const one = b.addSystemCommand(&[_][]const u8{ "ninja", "-C", build_dir, t });
        one.step.dependOn(&configure.step);
        const step = b.step(t, b.fmt("build only {s}", .{t}));
        step.dependOn(&one.step);
    }
}
