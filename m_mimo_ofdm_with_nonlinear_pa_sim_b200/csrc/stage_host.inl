// stage-level C ABI (to be filled in)
