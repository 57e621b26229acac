// swift-tools-version:5.9
// SwiftPM package of the B200 backend: a C module over libpiper_b200.so (include/piper_b200.h) and
// PiperCUDA, whose PiperCUDARuntime keeps the PiperMetalRuntime API
// (ocrickard/piper-swift Sources/PiperMetal/PiperMetalRuntime.swift:10-122).
// NOT BUILT in this repository's image (no Swift toolchain); see INTEGRATION.md.
import PackageDescription

// PiperCore / PiperONNX are the reference's own modules (audio types, PiperConfig, ONNXModel): either add the two
// targets below to the reference's Package.swift, or build this package beside a checkout of it.
let package = Package(
    name: "PiperCUDA",
    products: [.library(name: "PiperCUDA", targets: ["PiperCUDA"])],
    dependencies: [.package(name: "PiperSwift", path: "../../piper-swift")],
    targets: [
        .systemLibrary(name: "CPiperB200", path: "Sources/CPiperB200"),
        .target(name: "PiperCUDA",
                dependencies: ["CPiperB200",
                               .product(name: "PiperCore", package: "PiperSwift"),
                               .product(name: "PiperONNX", package: "PiperSwift")],
                path: "Sources/PiperCUDA"),
    ]
)
