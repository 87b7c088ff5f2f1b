"""Host-side harnesses that drive the device through the reference's own runtime C API (librt_capi ->
librtd_b200sv.so): used by bench.py's end-to-end leg and by the plugin tests."""
