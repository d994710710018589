// radix_host.inl — host side of the radix-partitioned layout (included inside b200join.cu's anonymous namespace).
void radix_build(H *, const void *, const uint8_t *, int, int64_t, float, int64_t) {
    throw Fail{B200J_ERR_UNSUPPORTED, "radix layout not built in"};
}
int64_t radix_probe(H *, const void *, const uint8_t *, int64_t, int64_t, int, bool, bool, int64_t, bool, ResultOwner *, int64_t *) {
    throw Fail{B200J_ERR_UNSUPPORTED, "radix layout not built in"};
}
void radix_release(H *) {}
