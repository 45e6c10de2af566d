{
  "targets": [{
    "target_name": "kmeans_napi",
    "sources": ["kmeans_napi.c"],
    "include_dirs": ["../include"],
    "libraries": ["-L<(module_root_dir)/../k-means-demo_b200", "-lkmeans", "-Wl,-rpath,<(module_root_dir)/../k-means-demo_b200"]
  }]
}
