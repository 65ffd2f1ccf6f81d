{
  "targets": [{
    "target_name": "tritonhash",
    "sources": ["tritonhash_napi.cc"],
    "include_dirs": ["../../include"],
    "libraries": ["-L<(module_root_dir)/../../downloader_b200", "-ltritonhash", "-Wl,-rpath,<(module_root_dir)/../../downloader_b200"],
    "cflags_cc": ["-std=c++17"],
    "defines": ["NAPI_VERSION=6"]
  }]
}
