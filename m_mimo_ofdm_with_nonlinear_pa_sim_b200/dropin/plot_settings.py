"""Plot styling helpers with the reference's names (plot_settings.py:8,50).  matplotlib is imported
lazily so that headless compute nodes without it can still import the simulation modules."""


def set_latex_plot_style(use_tex: bool = False, fig_width_in: float = 7.0, fig_height_in: float = None) -> None:
    import matplotlib as mpl
    if fig_height_in is None:
        fig_height_in = fig_width_in * (5 ** 0.5 - 1) / 2  # golden ratio
    mpl.rcParams.update({
        "text.usetex": bool(use_tex), "font.family": "serif", "font.size": 8, "axes.labelsize": 8,
        "legend.fontsize": 7, "xtick.labelsize": 7, "ytick.labelsize": 7,
        "figure.figsize": (fig_width_in, fig_height_in), "lines.linewidth": 1.0, "lines.markersize": 3,
    })


def reset_color_cycle():
    import matplotlib.pyplot as plt
    plt.gca().set_prop_cycle(None)
