{
  "targets": [{
    "target_name": "dusp_b200_addon",
    "sources": ["dusp_b200_addon.c"],
    "libraries": ["-L<(module_root_dir)/../../dusp_b200", "-ldusp_b200", "-Wl,-rpath,<(module_root_dir)/../../dusp_b200"]
  }]
}
