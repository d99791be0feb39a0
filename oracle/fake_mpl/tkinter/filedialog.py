def askdirectory(*a, **k):
    return ""


def askopenfilename(*a, **k):
    return ""
