def start_matlab(*a, **k):
    raise RuntimeError("MATLAB engine is not available (QuaDRiGa / 5G-Toolbox paths are out of scope)")
