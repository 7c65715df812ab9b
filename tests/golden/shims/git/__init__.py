class Repo:
    pass
