// gather_host.inl — host side of b200j_gather (included inside b200join.cu's anonymous namespace).
void do_gather(H *, const b200j_result *, const b200j_batch *, uint32_t, b200j_gathered *) {
    throw Fail{B200J_ERR_UNSUPPORTED, "gather not built in"};
}
void do_gathered_release(H *, b200j_gathered *) {}
